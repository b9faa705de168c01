"""Summarise an .ncu-rep here (no GPU): key raw metrics + top stall lines.  usage: ncu_top.py REPORT [N]"""
import csv, subprocess, sys, io
rep = sys.argv[1]; N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, u = rows[0], rows[1]
keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "lts__t_sectors.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_red.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_atom.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors_srcunit_tex_op_read.sum",
        "lts__t_sectors_srcunit_tex_op_read_lookup_miss.sum", "lts__t_sectors_srcunit_tex_op_read_lookup_hit.sum",
        "smsp__average_warp_latency_per_inst_issued.ratio", "lts__t_sectors_op_red.sum", "lts__t_sectors_op_atom.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors.avg.pct_of_peak_sustained_elapsed"]
for r in rows[2:]:
    print("== kernel", r[h.index("Kernel Name")] if "Kernel Name" in h else "")
    for a, b, c in zip(h, u, r):
        if a in keys or "issue_stalled" in a and "per_issue_active" in a and float(c or 0) > 0.5:
            print(f"  {a} [{b}] = {c}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; idx = {n: i for i, n in enumerate(h)}
data = [r for r in rows[hi + 1:] if len(r) == len(h)]
tot = sum(int(r[idx["# Samples"]]) for r in data)
print("total samples", tot)
order = sorted(range(len(data)), key=lambda i: -int(data[i][idx["# Samples"]]))[:N]
for i in order:
    r = data[i]
    print(f"  #{i:5d} {int(r[idx['# Samples']]):7d} {100*int(r[idx['# Samples']])/tot:5.1f}%  {r[idx['Source']].strip()[:60]:60s} lsb={r[idx['stall_long_sb']]} bar={r[idx['stall_barrier']]} lg={r[idx['stall_lg']]} mio={r[idx['stall_mio']]} ssb={r[idx['stall_short_sb']]} exec={r[idx['Instructions Executed']]}")
