"""Recipe for oracle/_ref: the UNMODIFIED reference modules of the hot path, taken from the reference checkout where
it lies (/root/reference, build container only) into oracle/_ref/ (git-ignored, travels to the GPU box with the
snapshot).  Nothing is edited; `bench.py --impl reference` and the cpu_baseline legs import it when present so the
timed CPU path is the reference's own code (kind "reference"), else they fall back to oracle/geo_oracle.py ("port").
TEST / MEASUREMENT INFRASTRUCTURE ONLY -- nothing under vqvae_b200/ may import it."""
import shutil
import sys
from pathlib import Path

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent / "_ref"
FILES = ["src/geo/__init__.py", "src/geo/geo_shortest_paths.py", "src/geo/knn_graph_optimized.py",
         "src/geo/kmeans_optimized.py", "src/geo/riemannian_metric.py"]


def main():
    if not REF.exists():
        print("make_ref: /root/reference not present; keeping the prebuilt oracle/_ref (if any)")
        return 0
    for rel in FILES:
        dst = OUT / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(REF / rel, dst)
    (OUT / "src" / "__init__.py").write_text("")  # package marker only (the reference's own src/__init__.py pulls in training code)
    (OUT / "PROVENANCE.txt").write_text("verbatim copies of " + ", ".join(FILES) + " from /root/reference (m4rch1n0/vqvae)\n")
    print(f"make_ref: {len(FILES)} reference modules -> {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
