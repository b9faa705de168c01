#!/usr/bin/env python
"""bench.py -- BASELINE.json metric: codebook build seconds & SSSP edge-relax/s (2.94M latents, K=512).

    python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
    python bench.py --impl reference ...                      (reference CPU implementation, bounded sample)

Workload (configs[1]): z = randn(60000,16,7,7; seed 42) -> 2.94M latents d=16, random-init FashionMNIST
spatial decoder (seed 0, BN train mode), k=20 union kNN, Riemannian edge lengths, LCC, K=512 kpp seed 42.
Setup runs the whole codebook build once on the GPU (reported under "build", seconds per stage); a timed
"step" is then one `dijkstra_multi_source`-shaped call: S=--sources-per-step independent SSSPs on that graph.
  value = ranks * S * nnz / t      directed-edge relaxations per second (each edge once per source)
  e2e   = same call from pinned HOST buffers (CSR upload + SSSP + D[S,n] float32 download in the timed region)
Inputs (the graph, 0.8 GB) exceed the 126 MB L2, so no explicit L2 flush is needed between steps.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FASHION_CFG = dict(in_channels=1, output_image_size=28, latent_dim=16, enc_channels=[64, 128, 256],
                   dec_channels=[256, 128, 64], recon_loss="mse", norm_type="batch", mse_use_sigmoid=True)
METRIC = "sssp_edge_relax_per_s"
UNIT = "edge-relax/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--images", type=int, default=60000, help="N of the (N,16,7,7) latent tensor")
    ap.add_argument("--grid", type=int, default=7)
    ap.add_argument("--sources-per-step", type=int, default=64)
    ap.add_argument("--slots", type=int, default=0, help="concurrent sources per launch (0 = auto)")
    ap.add_argument("--K", type=int, default=512)
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--ref-latents", type=int, default=100000, help="reference arm: size of the latent subsample")
    return ap.parse_args()


# ------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.rows, self.proc, self.idx = [], None, str(gpu_index)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", self.idx], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[1]) for r in self.rows if len(r) >= 8 and r[1].replace(".", "").isdigit()]
        mx = [float(r[2]) for r in self.rows if len(r) >= 8 and r[2].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 8 for n, v in zip(names, r[4:8]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------ reference arm
_CPU_W = None  # graph shared with forked workers (copy-on-write, no pickling)


def _dijkstra_chunk(src):
    from scipy.sparse.csgraph import dijkstra
    D = dijkstra(csgraph=_CPU_W, directed=False, indices=src)
    return float(np.asarray(D, dtype=np.float32)[:, ::4096].sum())


class CpuDijkstra:
    """Reference SSSP (scipy Dijkstra, as geo_shortest_paths.py:36-49 calls it) over a pool of `workers` forked
    processes that lives across steps: only the searches are timed, not process start-up."""

    def __init__(self, W, workers):
        import multiprocessing as mp
        global _CPU_W
        _CPU_W = W
        self.W, self.workers = W, workers
        self.pool = mp.get_context("fork").Pool(workers) if workers > 1 else None
        if self.pool is not None:
            self.pool.map(_dijkstra_chunk, [np.zeros(0, dtype=np.int64)] * workers)  # start the workers

    def rate(self, sources):
        chunks = [c for c in np.array_split(np.asarray(sources), self.workers) if len(c)]
        t0 = time.perf_counter()
        if self.pool is None:
            _dijkstra_chunk(chunks[0])
        else:
            self.pool.map(_dijkstra_chunk, chunks, chunksize=1)
        dt = time.perf_counter() - t0
        return len(sources) * self.W.nnz / dt, dt

    def close(self):
        if self.pool is not None:
            self.pool.close()
            self.pool.join()


def cpu_dijkstra_rate(W, sources, workers):
    cd = CpuDijkstra(W, workers)
    try:
        return cd.rate(sources)
    finally:
        cd.close()


def run_reference(a):
    """The reference's own CPU path (sklearn kNN + scipy Dijkstra through oracle/geo_oracle.py, which calls the
    same library entry points as src/geo/*) on a bounded sample; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    sys.path.insert(0, str(ROOT / "oracle"))
    import torch
    import geo_oracle as oracle
    cores = len(os.sched_getaffinity(0))
    M = a.ref_latents
    z = torch.randn(a.images, 16, a.grid, a.grid, generator=torch.Generator().manual_seed(42))
    zf = z.permute(0, 2, 3, 1).reshape(-1, 16)[:M].numpy()
    t0 = time.perf_counter()
    W, _ = oracle.knn_graph(zf, k=20, mode="distance", sym="union")  # Euclidean weights: CPU JVPs would take hours
    t_knn = time.perf_counter() - t0
    W = oracle.check_graph(W)
    workers = max(1, min(cores, 64))
    per_step = 4 * workers  # four sources per worker per step: a step is a few seconds of pure search time
    rng = np.random.RandomState(0)
    times = []
    cd = CpuDijkstra(W, workers)
    for i in range(a.warmup + a.steps):
        src = rng.choice(M, per_step, replace=False)
        rate, dt = cd.rate(src)
        if i >= a.warmup:
            times.append(dt)
    cd.close()
    t = float(np.sum(times))
    value = a.steps * per_step * W.nnz / t
    sample = (f"{per_step} sources/step on the union kNN graph (k=20, Euclidean weights) of the first {M} of "
              f"{a.images * a.grid * a.grid} latents, n={M}, nnz={W.nnz}; sklearn kNN took {t_knn:.1f}s (untimed)")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * t / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "FashionMNIST-shape geodesic codebook: 60k x 7 x 7 = 2.94M latents d=16, k=20 union, K=512, seed 42",
                       "sources_per_step": per_step, "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------ B200 arm
def run_b200(a):
    import torch
    import torch.distributed as dist
    from vqvae_b200 import _lib, device_graph as dg
    from vqvae_b200.models import SpatialVAE
    from vqvae_b200.pipeline import build_codebook_device

    # keep stdout to the one JSON line: libraries (NCCL's version banner) print to fd 1, so fd 1 is pointed at
    # stderr for the run and the line is written to the saved descriptor at the end
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    _lib.load()

    # ---- setup: the whole codebook build, once (this IS the first half of the metric) ----
    z4 = torch.randn(a.images, 16, a.grid, a.grid, generator=torch.Generator().manual_seed(42))
    torch.manual_seed(0)
    decoder = SpatialVAE(**FASHION_CFG).decoder.to(dev)
    if world > 1:  # NCCL connects lazily per protocol: take that out of the build seconds (process start-up, not work)
        for nel in (1 << 10, 1 << 24):
            w_in = torch.zeros(nel, dtype=torch.float32, device=dev)
            dist.all_gather([torch.empty_like(w_in) for _ in range(world)], w_in)
        del w_in
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = build_codebook_device(z4, decoder, k=20, sym="union", K=a.K, init="kpp", seed=42, batch_size=512,
                                verbose=False)
    torch.cuda.synchronize()
    build_total = time.perf_counter() - t0
    g = res.g_lcc
    n, nnz = g.n, g.nnz
    del z4
    S = a.sources_per_step
    slots = a.slots if a.slots > 0 else dg.default_slots(n)
    slots = max(1, min(slots, S, 32))
    row_ptr, adj, mean_w = g.adjacency(False)
    rng = np.random.RandomState(1234 + rank)
    D = torch.empty((S, n), dtype=torch.float32, device=dev)
    launches_per_step = -(-S // slots)  # lock-step batches of `slots` sources, one persistent launch each

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_sources(i):
        # medoids first (the K-source assignment call of kmeans_optimized.py:97), then seeded random vertices
        lo = (i * S) % max(1, len(res.medoids))
        src = np.asarray(res.medoids[lo:lo + S], dtype=np.int64)
        if len(src) < S:
            src = np.r_[src, rng.choice(n, S - len(src), replace=False)]
        return src

    # ---- device-resident timing ----
    for i in range(a.warmup):
        dg.sssp_multi(g, step_sources(i), slots=slots, out=D)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats_tot = dict(relaxations=0, rounds=0, scans=0)
    torch.cuda.cudart().cudaProfilerStart()  # `ncu --profile-from-start off` then lists exactly the timed launches
    ev0.record()
    for i in range(a.steps):
        _, st = dg.sssp_multi(g, step_sources(a.warmup + i), slots=slots, out=D, stats=True)
        for k_ in stats_tot:
            stats_tot[k_] += st[k_]
    ev1.record()
    barrier()
    torch.cuda.cudart().cudaProfilerStop()
    clocks = sampler.stop()
    ms = ev0.elapsed_time(ev1)
    t_all = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
    ms = float(t_all.item())
    value = world * a.steps * S * nnz / (ms * 1e-3)

    # ---- end-to-end through host buffers (pinned): upload CSR, SSSP, download D ----
    e2e = None
    if not a.skip_e2e:
        h_indptr = g.indptr.cpu().pin_memory()
        h_indices = g.indices.cpu().pin_memory()
        h_data = g.data.cpu().pin_memory()
        h_D = torch.empty((S, n), dtype=torch.float32).pin_memory()

        def e2e_step(i):
            gg = dg.DeviceCSR(h_indptr.to(dev, non_blocking=True), h_indices.to(dev, non_blocking=True),
                              h_data.to(dev, non_blocking=True), n)
            dg.sssp_multi_to_host(gg, step_sources(i), h_D, slots=slots)  # D2H of batch i overlaps batch i+1
            torch.cuda.synchronize()
            return float(h_D[0, 0])

        for i in range(min(a.warmup, 2)):
            e2e_step(i)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.steps):
            e2e_step(a.warmup + i)
        e1.record()
        barrier()
        t2 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        e2e = {"value": world * a.steps * S * nnz / (float(t2.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": int(h_indptr.numel() * 8 + h_indices.numel() * 4 + h_data.numel() * 4 + S * 4),
               "d2h_bytes_per_step": int(S * n * 4), "ms_per_step": float(t2.item()) / a.steps}
        del h_D

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    # ---- roofline of the dominant kernel (sssp_kernel): algorithmic bytes per launch / mean launch time ----
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    bytes_per_launch = (S / launches_per_step) * 16.0 * (nnz + n)  # 16 B per directed edge + 16 B per vertex, per source
    n_launch = a.steps * launches_per_step
    launch_ms = ms / n_launch
    achieved = bytes_per_launch / (launch_ms * 1e-3) / 1e9
    traffic = None
    try:
        traffic = json.loads((ROOT / "profiles" / "sssp_traffic.json").read_text()).get("dram_bytes_per_launch")
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                "kernel": "sssp_kernel", "launch_ms": launch_ms, "algorithmic_bytes_per_launch": bytes_per_launch,
                "relaxations_per_algorithmic_edge": stats_tot["relaxations"] / (a.steps * S * nnz)}
    # ---- CPU baseline: scipy Dijkstra on the very same graph, bounded sample ----
    cpu = None
    if not a.skip_cpu_baseline and world == 1:
        sys.path.insert(0, str(ROOT / "oracle"))
        import geo_oracle as oracle
        W = oracle.check_graph(g.to_scipy())
        cores = len(os.sched_getaffinity(0))
        # all host cores the box gives us, one source each (~10 s); every worker needs ~2 GB for scipy's fp64 copy
        workers = max(1, min(cores, 32))
        try:
            import psutil
            workers = max(1, min(workers, int(psutil.virtual_memory().available // (3 << 30))))
        except Exception:
            workers = min(workers, 4)
        src = step_sources(0)[:workers]
        rate, dt = cpu_dijkstra_rate(W, src, workers)
        # parity spot check on the timed sample's first source
        d_ref = oracle.sssp_multi(W, src[:1])
        d_gpu = dg.sssp_multi(g, src[:1]).cpu().numpy()
        cpu = {"value": rate, "unit": UNIT, "cores": workers, "kind": "port",
               "sample": f"{len(src)} sources of the same graph (n={n}, nnz={nnz}), one scipy Dijkstra per process, {dt:.1f}s",
               "host_cores_available": cores, "bit_equal_to_gpu": bool(np.array_equal(d_ref, d_gpu))}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "FashionMNIST-shape geodesic codebook: 60k x 7 x 7 = 2.94M latents d=16, k=20 union, K=512, seed 42",
                       "latents": int(res.z_flat.shape[0]), "n": n, "nnz": nnz, "sources_per_step": S, "slots": slots,
                       "delta": st["delta"], "mean_edge_weight": mean_w, "l2": "inputs larger than L2 (graph 0.8 GB)", "parallelism": f"sources x{world}"},
            "build": {"total_s": build_total, **{k_: round(v, 4) for k_, v in res.timings.items()}, "qe": res.qe,
                      "medoids_head": [int(x) for x in res.medoids[:5]]},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": n_launch, "clocks": clocks,
            "sssp_stats": {k_: int(v) for k_, v in stats_tot.items()}}
    sys.stdout.flush()
    os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
