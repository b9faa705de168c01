"""`build_codebook` CLI -- drop-in for the reference's src/scripts/build_codebook.py.

Same 17 flags and defaults (:109-131), same inputs (`z.pt` 4-D tensor, checkpoint dict with
'model_state_dict') and the same three artifacts in --out_dir:
  knn_graph_geodesic.npz  scipy CSR float32 of the LCC-restricted geodesic graph   (:65)
  codebook.pt             {'medoid_indices' int32 (K,), 'z_medoid' f32 (K,C), 'config' {...}}  (:79-102)
  codes.npy               int32 (N,H,W), -1 outside the LCC                         (:74-76,103)
All compute runs on the GPU through vqvae_b200.pipeline; there is no CPU fallback.
Launch under torchrun to shard the kNN and edge-length stages over several GPUs
(rank 0 writes the artifacts).
"""
import argparse
import os
from pathlib import Path

import numpy as np
import torch
from scipy import sparse

from vqvae_b200.models import SpatialVAE
from vqvae_b200.pipeline import build_codebook_device

CONFIG_KEYS = ["latents_path", "out_dir", "vae_ckpt_path", "in_channels", "output_image_size", "latent_dim",
               "enc_channels", "dec_channels", "recon_loss", "norm_type", "mse_use_sigmoid", "k", "sym", "K", "init",
               "seed", "batch_size"]


def _maybe_init_distributed():
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        if not dist.is_initialized():
            dist.init_process_group("nccl")
        return dist.get_rank()
    return 0


def main(args):
    """Builds a spatial codebook using a geodesic metric and saves artifacts."""
    if not torch.cuda.is_available():
        raise RuntimeError("build_codebook needs a CUDA (sm_100a) device: no CPU fallback exists")
    rank = _maybe_init_distributed()
    out_dir = Path(args.out_dir)
    out_dir.mkdir(parents=True, exist_ok=True)
    device = torch.device("cuda", torch.cuda.current_device())

    vae = SpatialVAE(in_channels=args.in_channels, output_image_size=args.output_image_size,
                     latent_dim=args.latent_dim, enc_channels=args.enc_channels, dec_channels=args.dec_channels,
                     recon_loss=args.recon_loss, norm_type=args.norm_type,
                     mse_use_sigmoid=args.mse_use_sigmoid).to(device)
    vae.load_state_dict(torch.load(args.vae_ckpt_path, map_location=device)["model_state_dict"])
    decoder = vae.decoder  # left in train mode exactly like the reference (:27-29): BN uses batch statistics

    z = torch.load(Path(args.latents_path), map_location="cpu").float()
    N, C, H, W = z.shape
    if rank == 0:
        print(f"Loaded spatial latents: N={N}, C={C}, H={H}, W={W}")
        print(f"Reshaped to: {(N * H * W, C)}")
        print(f"Building k-NN graph: N={N * H * W}, k={args.k}, method=cuda")

    res = build_codebook_device(z, decoder, k=args.k, sym=args.sym, K=args.K, init=args.init, seed=args.seed,
                                batch_size=args.batch_size, verbose=(rank == 0))
    if rank != 0:
        return res

    sparse.save_npz(out_dir / "knn_graph_geodesic.npz", res.g_lcc.to_scipy())
    z_lcc = res.z_flat[res.mask.bool()]
    z_medoid = z_lcc[torch.as_tensor(res.medoids, device=z_lcc.device)].cpu()
    codebook = {
        "medoid_indices": res.medoids.astype(np.int32),
        "z_medoid": z_medoid.float(),
        "config": {key: getattr(args, key) for key in CONFIG_KEYS},
    }
    torch.save(codebook, out_dir / "codebook.pt")
    np.save(out_dir / "codes.npy", res.codes)
    print(f"Quantization error: {res.qe:.3f}")
    print(f"Saved artifacts to: {out_dir}")
    print("Stage seconds: " + ", ".join(f"{k}={v:.3f}" for k, v in res.timings.items()))
    return res


def make_parser():
    p = argparse.ArgumentParser(description="Build a geodesic spatial codebook.")
    p.add_argument("--latents_path", type=str, required=True)
    p.add_argument("--out_dir", type=str, required=True)
    p.add_argument("--vae_ckpt_path", type=str, required=True)
    p.add_argument("--in_channels", type=int, required=True)
    p.add_argument("--output_image_size", type=int, required=True)
    p.add_argument("--latent_dim", type=int, required=True)
    p.add_argument("--enc_channels", type=int, nargs="+", required=True)
    p.add_argument("--dec_channels", type=int, nargs="+", required=True)
    p.add_argument("--recon_loss", type=str, required=True)
    p.add_argument("--norm_type", type=str, required=True)
    p.add_argument("--mse_use_sigmoid", action="store_true")
    p.add_argument("--k", type=int, default=20)
    p.add_argument("--sym", type=str, default="union")
    p.add_argument("--K", type=int, default=512)
    p.add_argument("--init", type=str, default="kpp")
    p.add_argument("--seed", type=int, default=42)
    p.add_argument("--batch_size", type=int, default=512)
    return p


if __name__ == "__main__":
    main(make_parser().parse_args())
