"""torchrun worker for tests/test_gpu_multi.py: the sharded build and the source-sharded K-medoids calls must give
the same bits as a single-GPU run.  Prints one JSON line from rank 0."""
import hashlib
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from conftest import c1_latents, fashion_decoder  # noqa: E402


def sha(*arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from vqvae_b200.pipeline import build_codebook_device
    from vqvae_b200.geo import kmeans_optimized as km
    import contextlib
    import io
    z4, _ = c1_latents()
    dec = fashion_decoder().cuda()
    out = {"world": world}
    with contextlib.redirect_stdout(io.StringIO()):
        sharded = build_codebook_device(z4, dec, k=20, sym="union", K=64, verbose=False)
        single = build_codebook_device(z4, dec, k=20, sym="union", K=64, verbose=False, distributed=False)
        out["build_equal"] = bool(np.array_equal(sharded.medoids, single.medoids) and np.array_equal(sharded.codes, single.codes)
                                  and sharded.qe == single.qe and torch.equal(sharded.g_lcc.data, single.g_lcc.data))
        out["digest"] = sha(sharded.medoids.astype(np.int64), sharded.codes.astype(np.int32))
        g = single.g_lcc
        # K independent sources split across ranks + allreduce-MIN of (dist32 bits, k) keys
        grp = dist.group.WORLD
        med_r, asg_r, qe_r = km.fit_kmedoids_optimized(g, K=96, init="random", seed=5, dist_group=grp)
        med_1, asg_1, qe_1 = km.fit_kmedoids_optimized(g, K=96, init="random", seed=5)
        out["random_init_equal"] = bool(np.array_equal(med_r, med_1) and np.array_equal(asg_r, asg_1) and qe_r == qe_1)
        # collective upload of a host matrix: 1/world per rank + NCCL all-gather == plain upload, and the public call on it
        from vqvae_b200.geo import geo_shortest_paths as gsp
        Wh = g.to_scipy()
        g_sh = gsp.graph_to_device(Wh, dist_group=grp)
        out["sharded_upload_equal"] = bool(torch.equal(g_sh.indptr, g.indptr) and torch.equal(g_sh.indices, g.indices)
                                           and torch.equal(g_sh.data, g.data))
        src = [3, 1000 + rank, 19999]
        out["sssp_on_sharded_equal"] = bool(np.array_equal(gsp.dijkstra_multi_source(g_sh, src), gsp.dijkstra_multi_source(Wh, src)))
        a_sh = km.assign_points_to_medoids(g, single.medoids, dist_group=grp)
        out["assign_equal"] = bool(np.array_equal(a_sh, single.assign))
        out["qe_equal"] = bool(km.compute_quantization_error(g, single.medoids, single.assign, dist_group=grp) == single.qe)
    flags = torch.tensor([int(all(v for k, v in out.items() if k.endswith("_equal")))], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    out["all_ranks_ok"] = bool(flags.item())
    if rank == 0:
        print("MULTI_GPU_RESULT " + json.dumps(out), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
