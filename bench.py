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
CIFAR_CFG = dict(in_channels=3, output_image_size=32, latent_dim=32, enc_channels=[64, 128, 256],
                 dec_channels=[256, 128, 64], recon_loss="mse", norm_type="batch", mse_use_sigmoid=False)


def workload(a):
    """--workload -> (name, latent tensor factory (N,C,H,W), decoder config, K).
    c2 (default, BASELINE configs[1], the headline): randn(60000,16,7,7) = 2.94M latents d=16, Fashion decoder.
    c3 (configs[2]): randn(50000,32,8,8) = 3.2M latents d=32, CIFAR decoder (JVP output 3x8x8).
    c5:M,d,K (configs[4]): randn(M,d,1,1) flat synthetic latents, Fashion-shape decoder with latent_dim=d."""
    import torch
    w = a.workload
    if w == "c2":
        n_img, g = a.images, a.grid
        name = f"FashionMNIST-shape geodesic codebook: {n_img//1000}k x {g} x {g} = {n_img*g*g/1e6:.2f}M latents d=16, k=20 union, K={a.K}, seed 42"
        return name, (lambda: torch.randn(n_img, 16, g, g, generator=torch.Generator().manual_seed(42))), FASHION_CFG, a.K
    if w == "c3":
        name = f"CIFAR-shape geodesic codebook: 50k x 8 x 8 = 3.2M latents d=32, 3x32x32 decoder, k=20 union, K={a.K}, seed 42"
        return name, (lambda: torch.randn(50000, 32, 8, 8, generator=torch.Generator().manual_seed(42))), CIFAR_CFG, a.K
    if w.startswith("c5:"):
        M, d, K = (int(float(x)) for x in w[3:].split(","))
        name = f"scaling sweep: {M/1e6:g}M synthetic latents d={d}, k=20 union, K={K}, seed 42"
        return name, (lambda: torch.randn(M, d, 1, 1, generator=torch.Generator().manual_seed(42))), dict(FASHION_CFG, latent_dim=d), K
    raise SystemExit(f"unknown --workload {w}")
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
    ap.add_argument("--workload", default="c2", help="c2 (headline) | c3 | c5:M,d,K")
    ap.add_argument("--kernel", default=None, help="SSSP kernel: slot | vm (default: library default)")
    ap.add_argument("--knn-slice", type=int, default=20000, help="cpu_baseline_build: sklearn kNN query rows")
    ap.add_argument("--jvp-batches", type=int, default=64, help="cpu_baseline_build: torch-CPU JVP 512-edge batches")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-e2e", action="store_true")
    ap.add_argument("--ref-latents", type=int, default=400000,
                    help="reference arm: size of the latent subsample (large enough that the graph leaves the CPU caches)")
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


_REF_DIR = ROOT / "oracle" / "_ref"  # unmodified reference modules (oracle/make_ref.py); absent -> oracle port


def _ref_geo():
    """The reference's own src.geo modules from oracle/_ref, or None."""
    if not (_REF_DIR / "src" / "geo" / "geo_shortest_paths.py").exists():
        return None
    if str(_REF_DIR) not in sys.path:
        sys.path.insert(0, str(_REF_DIR))
    import importlib
    return importlib.import_module("src.geo.geo_shortest_paths"), importlib.import_module("src.geo.knn_graph_optimized")


def _dijkstra_chunk(src):
    if len(src) == 0:
        return 0.0
    ref = _ref_geo() if _CPU_USE_REF else None
    if ref is not None:  # the reference's function itself (geo_shortest_paths.py:24-50), float32 result
        D = ref[0].dijkstra_multi_source(_CPU_W, np.asarray(src))
    else:
        from scipy.sparse.csgraph import dijkstra
        D = np.asarray(dijkstra(csgraph=_CPU_W, directed=False, indices=src), dtype=np.float32)
    return float(D[:, ::4096].sum())


_CPU_USE_REF = False


def _dijkstra_rows(src):
    """Same search, rows kept (float32, as geo_shortest_paths.py:50 returns them) for the parity check."""
    from scipy.sparse.csgraph import dijkstra
    if len(src) == 0:
        return np.zeros((0, _CPU_W.shape[0]), np.float32)
    return np.asarray(dijkstra(csgraph=_CPU_W, directed=False, indices=src), dtype=np.float32)


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

    def rate(self, sources, keep_rows=False):
        chunks = [c for c in np.array_split(np.asarray(sources), self.workers) if len(c)]
        fn = _dijkstra_rows if keep_rows else _dijkstra_chunk
        t0 = time.perf_counter()
        out = [fn(chunks[0])] if self.pool is None else self.pool.map(fn, chunks, chunksize=1)
        dt = time.perf_counter() - t0
        if keep_rows:
            return len(sources) * self.W.nnz / dt, dt, np.concatenate(out, 0)
        return len(sources) * self.W.nnz / dt, dt

    def close(self):
        if self.pool is not None:
            self.pool.close()
            self.pool.join()


def cpu_dijkstra_rate(W, sources, workers, keep_rows=False):
    cd = CpuDijkstra(W, workers)
    try:
        return cd.rate(sources, keep_rows)
    finally:
        cd.close()


def cpu_build_baselines(a, res, dec_cfg, oracle, dijkstra_s_per_source):
    """Reference CPU stages of the BUILD on bounded samples of the same inputs, each
    extrapolated linearly and checked against the GPU result on the sample:
      kNN      sklearn NearestNeighbors (knn_graph_optimized.py:39-42): `knn_slice` query rows against the full set;
      JVP      torch-CPU functional.jvp edge lengths (riemannian_metric.py:37-66): the first `jvp_batches` batches;
      Dijkstra the reference runs (K-1)+K+K single-threaded scipy searches (kmeans_optimized.py:40-44,97,125)."""
    import torch
    from sklearn.neighbors import NearestNeighbors
    from vqvae_b200 import device_graph as dg
    from vqvae_b200.models import SpatialVAE
    zf = res.z_flat.cpu().numpy()
    M = zf.shape[0]
    q = min(a.knn_slice, M)
    t0 = time.perf_counter()
    nn_ = NearestNeighbors(n_neighbors=21, algorithm="auto", metric="euclidean").fit(zf)
    _, idx_cpu = nn_.kneighbors(zf[:q], return_distance=True)
    t_knn = time.perf_counter() - t0
    idx_gpu, _, _ = dg.knn_topk(res.z_flat, 21, 0, q)
    same_rows = int((idx_gpu.cpu().numpy() == idx_cpu).all(axis=1).sum())
    nb_total = -(-res.ei.numel() // 512)
    nb = min(a.jvp_batches, nb_total)
    ne = min(nb * 512, res.ei.numel())
    ei, ej = res.ei[:ne].cpu().numpy(), res.ej[:ne].cpu().numpy()
    torch.manual_seed(0)
    dec_cpu = SpatialVAE(**dec_cfg).decoder  # same seed -> same weights as the GPU decoder; train mode
    zs, ze = torch.from_numpy(zf[ei]), torch.from_numpy(zf[ej])
    t0 = time.perf_counter()
    L_cpu = oracle.edge_lengths(dec_cpu, zs, ze, 512).numpy()
    t_jvp = time.perf_counter() - t0
    L_gpu = res.edge_len[:ne].cpu().numpy()
    rel = np.abs(L_gpu - L_cpu) / np.maximum(np.abs(L_cpu), 1e-30)
    K = len(res.medoids)
    n_dij = (K - 1) + K + K
    out = {"knn": {"seconds": t_knn, "sample": f"{q} of {M} query rows against all {M} latents", "fraction": q / M,
                   "extrapolated_s": t_knn * M / q, "index_rows_identical_to_gpu": f"{same_rows}/{q}"},
           "edge_len": {"seconds": t_jvp, "sample": f"first {nb} of {nb_total} 512-edge batches", "fraction": nb / nb_total,
                        "extrapolated_s": t_jvp * nb_total / nb, "edges_beyond_1e-5_vs_gpu": int((rel > 1e-5).sum()),
                        "edges": int(ne), "median_rel": float(np.median(rel)), "max_rel": float(rel.max())},
           "dijkstra": {"seconds_per_source": dijkstra_s_per_source, "sources_in_reference_build": n_dij,
                        "extrapolated_s": dijkstra_s_per_source * n_dij, "threads": 1,
                        "note": "scipy's Dijkstra is single-threaded and the reference calls it serially"},
           "threads": {"sklearn/torch": "library defaults (all cores)", "cpu_count": os.cpu_count(),
                       "affinity": len(os.sched_getaffinity(0))}}
    out["total_extrapolated_s"] = out["knn"]["extrapolated_s"] + out["edge_len"]["extrapolated_s"] + out["dijkstra"]["extrapolated_s"]
    return out


def run_reference(a):
    """The reference's own CPU path (sklearn kNN + scipy Dijkstra through oracle/geo_oracle.py, which calls the
    same library entry points as src/geo/*) on a bounded sample; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    global _CPU_USE_REF
    sys.path.insert(0, str(ROOT / "oracle"))
    import torch
    ref = _ref_geo()
    _CPU_USE_REF = ref is not None
    kind = "reference" if ref is not None else "port"
    cores = len(os.sched_getaffinity(0))
    try:  # torchrun exports OMP_NUM_THREADS=1: give sklearn's brute kNN the host cores back
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=cores)
    except Exception:
        pass
    M = a.ref_latents
    z = torch.randn(a.images, 16, a.grid, a.grid, generator=torch.Generator().manual_seed(42))
    zf = z.permute(0, 2, 3, 1).reshape(-1, 16)[:M].numpy()
    t0 = time.perf_counter()
    if ref is not None:
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):  # Euclidean weights: CPU JVPs of 2.94M latents would take hours
            W, _ = ref[1].build_knn_graph_auto(zf, k=20, metric="euclidean", mode="distance", sym="union")
        W = ref[0].ensure_valid_graph(W)
    else:
        import geo_oracle as oracle
        W, _ = oracle.knn_graph(zf, k=20, mode="distance", sym="union")
        W = oracle.check_graph(W)
    t_knn = time.perf_counter() - t0
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
            "config": {"workload": workload(a)[0], "sources_per_step": per_step, "sample": sample},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": kind, "sample": sample,
                             "code": "oracle/_ref/src/geo (unmodified reference modules)" if kind == "reference" else "oracle/geo_oracle.py"},
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
    wname, make_z, dec_cfg, K = workload(a)
    z4 = make_z()
    torch.manual_seed(0)
    decoder = SpatialVAE(**dec_cfg).decoder.to(dev)
    if world > 1:  # NCCL connects lazily per protocol: take that out of the build seconds (process start-up, not work)
        for nel in (1 << 10, 1 << 24):
            w_in = torch.zeros(nel, dtype=torch.float32, device=dev)
            dist.all_gather([torch.empty_like(w_in) for _ in range(world)], w_in)
        del w_in
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    res = build_codebook_device(z4, decoder, k=20, sym="union", K=K, init="kpp", seed=42, batch_size=512,
                                verbose=False)
    torch.cuda.synchronize()
    build_total = time.perf_counter() - t0
    g = res.g_lcc
    n, nnz = g.n, g.nnz
    M_lat, d_lat = int(res.z_flat.shape[0]), int(res.z_flat.shape[1])
    E_edges = int(res.ei.numel())
    import hashlib
    digest = hashlib.sha256(np.ascontiguousarray(np.asarray(res.medoids, dtype=np.int64)).tobytes()
                            + np.ascontiguousarray(res.codes.astype(np.int32)).tobytes()).hexdigest()
    S = a.sources_per_step
    slots = a.slots if a.slots > 0 else dg.default_slots(n, S)
    slots = max(1, min(slots, S, 32))
    kernel = dg._pick_kernel(a.kernel, S, False)
    if kernel in ("vm", "vm32"):
        slots = min(S, 32 if kernel == "vm32" else _lib.load().geocb_sssp_vm_slots())
    row_ptr, adj, mean_w = g.adjacency(False)
    rng = np.random.RandomState(1234 + rank)
    D = torch.empty((S, n), dtype=torch.float32, device=dev)
    # slot-major kernel: ONE persistent launch per step, the S sources stream through `slots` live tables;
    # vertex-major kernels: one launch per `slots` sources
    launches_per_step = 1 if kernel == "slot" else -(-S // slots)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_sources(i):
        # the K-source assignment call of kmeans_optimized.py:97 split across ranks: every rank takes its own
        # S consecutive medoids of step i (disjoint across ranks), topped up with seeded random vertices
        med = np.asarray(res.medoids, dtype=np.int64)
        lo = ((i * world + rank) * S) % max(1, len(med))
        src = med[lo:lo + S]
        if len(src) < S:
            src = np.r_[src, rng.choice(n, S - len(src), replace=False)]
        return src

    # ---- device-resident timing ----
    for i in range(a.warmup):
        dg.sssp_multi(g, step_sources(i), slots=slots, out=D, kernel=kernel)
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats_tot = dict(relaxations=0, rounds=0, scans=0)
    torch.cuda.cudart().cudaProfilerStart()  # `ncu --profile-from-start off` then lists exactly the timed launches
    ev0.record()
    for i in range(a.steps):
        _, st = dg.sssp_multi(g, step_sources(a.warmup + i), slots=slots, out=D, stats=True, kernel=kernel)
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

    # ---- multi-GPU product path: the K-source assignment, sources split across ranks + allreduce-MIN of keys ----
    assign_sharded = None
    if world > 1:
        import contextlib
        import io
        from vqvae_b200.geo import kmeans_optimized as km
        with contextlib.redirect_stdout(io.StringIO()):
            km.assign_points_to_medoids(g, res.medoids, dist_group=dist.group.WORLD)  # warm-up
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            a_sh = km.assign_points_to_medoids(g, res.medoids, dist_group=dist.group.WORLD)
            s1.record()
            barrier()
        t3 = torch.tensor([s0.elapsed_time(s1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t3, op=dist.ReduceOp.MAX)
        assign_sharded = {"K": int(len(res.medoids)), "ms": float(t3.item()), "equal_to_replicated": bool(np.array_equal(a_sh, res.assign)),
                          "collective": f"allreduce-MIN of {n} int64 (f32 bits<<32|k) keys"}

    # ---- end-to-end through the PUBLIC call the reference's callers bind: dijkstra_multi_source(W_scipy, sources)
    #      (src/geo/geo_shortest_paths.py:24-50): host CSR in, host (S,n) float32 out, every copy inside the timed
    #      region.  N>1: the upload is split across the ranks (1/N of the CSR per PCIe link + NCCL all-gather over
    #      NVLink); every rank then runs its own sources and downloads its own rows. ----
    e2e = e2e_dev = None
    if not a.skip_e2e:
        from vqvae_b200.geo import geo_shortest_paths as gsp
        W_host = g.to_scipy()  # every rank holds the host matrix (the caller's data), as in an SPMD launch

        def e2e_step(i):
            src = step_sources(i)
            if world == 1:
                Dh = gsp.dijkstra_multi_source(W_host, src)
                return float(Dh[0, 0])
            # each rank uploads 1/N of the CSR over its own PCIe link, NCCL all-gather over NVLink, validation and
            # symmetry check on every rank; then every rank searches its own sources and downloads its own rows
            gg = gsp.graph_to_device(W_host, dist_group=dist.group.WORLD)
            Dh = gsp.dijkstra_multi_source(gg, src)
            return float(Dh[0, 0])

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
        csr_bytes = int((n + 1) * 8 + nnz * 4 + nnz * 4)
        e2e = {"value": world * a.steps * S * nnz / (float(t2.item()) * 1e-3), "unit": UNIT,
               "h2d_bytes_per_step": csr_bytes + S * 8, "d2h_bytes_per_step": int(world * S * n * 4),
               "ms_per_step": float(t2.item()) / a.steps,
               "api": "vqvae_b200.geo.dijkstra_multi_source(scipy CSR, sources) -> numpy (S,n) float32"
                      + ("" if world == 1 else "; graph_to_device(W, dist_group): 1/N of the CSR uploaded per rank + NCCL all-gather, per-rank sources")}

        # previous rounds' figure, one layer below the public API: pinned CSR -> DeviceCSR -> pinned D
        h_indptr = g.indptr.cpu().pin_memory()
        h_indices = g.indices.cpu().pin_memory()
        h_data = g.data.cpu().pin_memory()
        h_D = torch.empty((S, n), dtype=torch.float32).pin_memory()

        def dev_step(i):
            gg = dg.DeviceCSR(h_indptr.to(dev, non_blocking=True), h_indices.to(dev, non_blocking=True),
                              h_data.to(dev, non_blocking=True), n)
            # (copy-out batches of 16 rows: the last batch's device->host copy is the exposed one)
            dg.sssp_multi_to_host(gg, step_sources(i), h_D, slots=a.slots if a.slots > 0 else None, kernel=kernel)
            torch.cuda.synchronize()

        dev_step(0)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.steps):
            dev_step(a.warmup + i)
        e1.record()
        barrier()
        t4 = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t4, op=dist.ReduceOp.MAX)
        e2e_dev = {"value": world * a.steps * S * nnz / (float(t4.item()) * 1e-3), "unit": UNIT,
                   "ms_per_step": float(t4.item()) / a.steps, "api": "device_graph.DeviceCSR + sssp_multi_to_host on pinned tensors"}
        del h_D, W_host

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
        if kernel == "slot" and a.workload == "c2":  # the committed ncu capture is of this kernel on this workload
            tj = json.loads((ROOT / "profiles" / "sssp_traffic.json").read_text())
            # (captured on a 16-source launch: DRAM bytes scale with the sources a launch carries)
            traffic = tj.get("dram_bytes_per_launch") * (S / launches_per_step) / tj.get("sources_per_launch", 16)
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                "kernel": {"vm": "sssp_vm_kernel", "vm32": "sssp_vm32_kernel"}.get(kernel, "sssp_kernel"), "launch_ms": launch_ms, "algorithmic_bytes_per_launch": bytes_per_launch,
                "relaxations_per_algorithmic_edge": stats_tot["relaxations"] / (a.steps * S * nnz)}
    # ---- CPU baseline: scipy Dijkstra on the very same graph, bounded sample ----
    cpu = cpu_build = None
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
        rate, dt, rows = cpu_dijkstra_rate(W, src, workers, keep_rows=True)
        # parity on EVERY source of the timed sample: float32 rows bit-equal to the GPU's
        d_gpu = dg.sssp_multi(g, src, kernel=kernel).cpu().numpy()
        n_equal = int(sum(np.array_equal(rows[i], d_gpu[i]) for i in range(len(src))))
        cpu = {"value": rate, "unit": UNIT, "cores": workers, "kind": "port",
               "sample": f"{len(src)} sources of the same graph (n={n}, nnz={nnz}), one scipy Dijkstra per process, {dt:.1f}s",
               "host_cores_available": cores, "bit_equal_sources": f"{n_equal}/{len(src)}",
               "bit_equal_to_gpu": bool(n_equal == len(src))}
        del rows, d_gpu
        try:
            cpu_build = cpu_build_baselines(a, res, dec_cfg, oracle, dt)  # dt = one search per process in parallel
        except Exception as ex:  # a baseline must never cost the bench line
            cpu_build = {"error": repr(ex)}
    # ---- per-stage rooflines of the build (reported next to the SSSP kernel's) ----
    tms = res.timings
    h1, h2 = dec_cfg["dec_channels"][1] * 4, dec_cfg["dec_channels"][2] * 16
    P_out = dec_cfg["in_channels"] * (64 if dec_cfg["output_image_size"] == 32 else 16)
    jvp_flops = 2.0 * (d_lat * h1 + h1 * h2 + h2 * P_out) * 4.0 * E_edges  # 2 end-points x (primal + tangent) rows per edge
    t_knn, t_edge = max(tms.get("knn", 0.0), 1e-9), max(tms.get("edge_len", 0.0), 1e-9)
    stage_rooflines = {
        "knn": {"bound": "tensor", "scores_per_s": M_lat * float(M_lat) / world / t_knn, "unit": "pairwise scores/s per GPU",
                "useful_flop_per_s": 2.0 * M_lat * float(M_lat) * d_lat / world / t_knn,
                "note": "select-bound, not tensor-bound: DESIGN.md 4.2 (scores/s vs the measured TMEM-read ceiling)"},
        "edge_len": {"bound": "tensor", "achieved": jvp_flops / world / t_edge / 1e12, "unit": "TFLOP/s per GPU (fp32-grade)",
                     "flop": jvp_flops, "note": "decoder JVP as three GEMMs per row; fp32 accuracy required (1e-5)"}}
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": wname, "latents": M_lat, "d": d_lat, "n": n, "nnz": nnz, "sources_per_step": S,
                       "slots": slots, "kernel": kernel, "delta": st["delta"], "mean_edge_weight": mean_w,
                       "l2": "inputs larger than L2 (graph 0.8 GB)",
                       "parallelism": f"sources x{world} (disjoint medoid sources per rank, no data-path collective)"},
            "build": {"total_s": build_total, **{k_: round(v, 4) for k_, v in res.timings.items()}, "qe": res.qe,
                      "medoids_head": [int(x) for x in res.medoids[:5]], "digest_medoids_codes_sha256": digest,
                      "edges": E_edges, "stage_rooflines": stage_rooflines},
            "roofline": roofline, "cpu_baseline": cpu, "cpu_baseline_build": cpu_build, "e2e": e2e,
            "e2e_device_graph": e2e_dev, "assign_sharded": assign_sharded, "gpu_launches": n_launch, "clocks": clocks,
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
