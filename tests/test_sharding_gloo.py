"""Multi-rank host logic on CPU (gloo, world_size 2): shard boundaries, the row all-gather used for kNN lists
and edge lengths, and the source split of the SSSP bench.  No CUDA needed."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vqvae_b200.sharding import all_gather_rows, shard_range, allreduce_min_assignment


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_shard_range_covers_and_respects_batches():
    for total, world, mult in [(0, 2, 1), (1, 2, 1), (10, 3, 1), (316157, 2, 512), (316157, 8, 512), (1000, 8, 512)]:
        pieces = [shard_range(total, r, world, mult) for r in range(world)]
        assert pieces[0][0] == 0 and pieces[-1][1] == total
        for (lo, hi), (lo2, _) in zip(pieces, pieces[1:]):
            assert hi == lo2 and lo <= hi
        for lo, hi in pieces:
            assert lo % mult == 0 or lo == total  # whole 512-edge BN batches per rank (reference batch order)
        sizes = [-(-(hi - lo) // mult) for lo, hi in pieces]
        assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, total, mult, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = torch.arange(total * 3, dtype=torch.float32).view(total, 3)
        lo, hi = shard_range(total, rank, world, mult)
        got = all_gather_rows(full[lo:hi].clone(), total, world, rank, None, dist, multiple=mult)
        ok = torch.equal(got, full)
        # 1-D payload (edge lengths) and an int payload (kNN index lists)
        v = torch.arange(total, dtype=torch.int32)
        got_v = all_gather_rows(v[lo:hi].clone(), total, world, rank, None, dist, multiple=mult)
        ok = ok and torch.equal(got_v, v)
        # SSSP assignment combine: lexicographic (dist32 bits, k) min over source shards == global argmin
        rng = np.random.RandomState(7)
        D = rng.rand(8, 50).astype(np.float32)
        D[:, 5] = 1.0  # ties -> lowest k must win
        klo, khi = shard_range(8, rank, world)
        key = (torch.from_numpy(D[klo:khi].view(np.uint32).astype(np.int64)) << 32) | \
            torch.arange(klo, khi, dtype=torch.int64)[:, None]
        best = key.min(dim=0).values
        dist.all_reduce(best, op=dist.ReduceOp.MIN)
        ok = ok and np.array_equal((best & 0xFFFFFFFF).numpy(), D.argmin(axis=0))
        # the product helper (vqvae_b200.sharding.allreduce_min_assignment, used by kmeans_optimized._labels_for):
        # per-rank Voronoi partials incl. unreachable vertices (+inf, -1), a vertex only one rank reaches, f32 ties
        D2 = rng.rand(8, 60).astype(np.float32)
        D2[:, 3] = np.inf          # nobody reaches vertex 3
        D2[:4, 9] = np.inf         # only the second shard reaches vertex 9
        D2[:, 11] = 0.25           # all tie -> k = 0
        D2[2, 12] = D2[6, 12] = 0.0  # cross-rank tie -> k = 2
        part = D2[klo:khi]
        dmin = torch.from_numpy(part.min(axis=0).copy())
        lab = torch.from_numpy((part.argmin(axis=0) + klo).astype(np.int32))
        lab[torch.isinf(dmin)] = -1
        allreduce_min_assignment(dmin, lab, dist)
        exp_lab = D2.argmin(axis=0).astype(np.int32)
        exp_lab[np.isinf(D2.min(axis=0))] = -1
        ok = ok and np.array_equal(lab.numpy(), exp_lab) and np.array_equal(dmin.numpy(), D2.min(axis=0))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("total,mult", [(1237, 1), (5000, 512), (300, 512)])
def test_all_gather_rows_world2(total, mult):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, mult, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok in res), res
