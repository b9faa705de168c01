import torch, time, os
torch.backends.cuda.matmul.allow_tf32 = False
a = torch.randn(131072, 512, device="cuda"); b = torch.randn(512, 1024, device="cuda")
ref = (a[:4096].double() @ b.double())
for name in ["fp32"]:
    for _ in range(3): c = a @ b
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): c = a @ b
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
    err = ((c[:4096].double() - ref).abs().max() / ref.abs().max()).item()
    print(os.environ.get("CUBLAS_EMULATE_SINGLE_PRECISION"), name, f"{dt*1e3:.3f} ms  {2*131072*512*1024/dt/1e12:.1f} TFLOP/s  relerr {err:.2e}", torch.version.cuda, torch.backends.cuda.preferred_blas_library())
torch.backends.cuda.matmul.allow_tf32 = True
for _ in range(3): c = a @ b
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(10): c = a @ b
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
err = ((c[:4096].double() - ref).abs().max() / ref.abs().max()).item()
print("tf32", f"{dt*1e3:.3f} ms  {2*131072*512*1024/dt/1e12:.1f} TFLOP/s  relerr {err:.2e}")
