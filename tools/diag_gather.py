"""Timing of the pipeline's two all-gathers and the scatter behind them (torchrun, N >= 2); analysis only."""
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from vqvae_b200.sharding import all_gather_rows, shard_range  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
M, E = 2_940_000, 45_709_818


def t():
    torch.cuda.synchronize()
    return time.perf_counter()


dist.barrier()
lo, hi = shard_range(M, rank, world)
idx = torch.randint(0, M, (hi - lo, 21), dtype=torch.int32, device=dev)
dst = torch.rand((hi - lo, 21), device=dev)
elo, ehi = shard_range(E, rank, world, multiple=512)
Lp = torch.rand(ehi - elo, device=dev)
pos = torch.randperm(2 * E, device=dev)[:E].to(torch.int32)
data = torch.zeros(2 * E, device=dev)
for rep in range(3):
    t0 = t(); a = all_gather_rows(idx, M, world, rank, None, dist); t1 = t()
    b = all_gather_rows(dst, M, world, rank, None, dist); t2 = t()
    L = all_gather_rows(Lp, E, world, rank, None, dist, multiple=512); t3 = t()
    data[pos] = L; t4 = t()
    data[pos.long()] = L; t5 = t()
    if rank == 0:
        print(f"rep {rep}: gather idx {1e3*(t1-t0):.1f} ms  dst {1e3*(t2-t1):.1f}  L {1e3*(t3-t2):.1f}  scatter(int32 idx) {1e3*(t4-t3):.1f}  scatter(int64 idx) {1e3*(t5-t4):.1f}", flush=True)
dist.destroy_process_group()
