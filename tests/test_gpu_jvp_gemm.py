"""tcgen05 3xTF32 GEMM of the decoder-JVP layers (csrc/jvp_gemm.cu) against a float64 product: fp32-grade accuracy
(the 1e-5 edge-length bar leaves no room for plain TF32's 2.7e-4)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("M,K,N", [(256, 16, 256), (512, 512, 1024), (4096, 512, 1024), (131072, 512, 1024), (768, 1024, 256)])
def test_gemm3x_matches_float64(M, K, N):
    from vqvae_b200 import device_graph as dg
    g = torch.Generator(device="cuda").manual_seed(M + K)
    a = torch.randn(M, K, generator=g, device="cuda")
    a[:, ::7] = 0.0  # post-ReLU activations are half zeros
    w = (torch.randn(K, N, generator=g, device="cuda") * 0.05).contiguous()
    c = dg.gemm3x(a, w)
    assert c is not None and c.shape == (M, N)
    rows = torch.arange(0, M, max(1, M // 512), device="cuda")
    ref = a[rows].double() @ w.double()
    scale = (a[rows].double().abs() @ w.double().abs())  # condition-aware bound: |c - ref| <= eps * sum |a||w|
    err = ((c[rows].double() - ref).abs() / scale.clamp_min(1e-30)).max().item()
    cub = ((a[rows] @ w).double() - ref).abs().div(scale.clamp_min(1e-30)).max().item()
    print(f"M={M} K={K} N={N}: gemm3x err {err:.3g} (cuBLAS fp32 {cub:.3g}) of sum|a||w|")
    assert err < 2.5e-6  # measured 1.3-1.9e-6 (fp32 accumulation in the tensor core truncates); plain TF32 would be ~5e-4
    assert dg.gemm3x(a[:100], w) is None  # unsupported shape -> caller keeps the library GEMM


@pytest.mark.parametrize("kin,C,HW", [(16, 128, 4), (32, 64, 8), (16, 512, 1)])
def test_fused_input_layer_matches_unfused(kin, C, HW):
    """geocb_lin_bn_relu_jvp_to_image (first linear map fused into the BatchNorm-JVP kernel) against the unfused pair
    SGEMM + geocb_bn_relu_jvp_to_image: same operand image up to the SGEMM's summation order (1e-6 relative)."""
    from vqvae_b200 import device_graph as dg
    g = torch.Generator().manual_seed(kin + C)
    n, groups, ld = 1024, 2, C * HW
    z = torch.randn(n, kin, generator=g).cuda()
    dz = (torch.randn(n, kin, generator=g) * 0.1).cuda()
    w = (torch.randn(kin, ld, generator=g) / kin ** 0.5).cuda()
    pb = torch.randn(ld, generator=g).cuda()
    bw, bb = (torch.rand(C, generator=g) + 0.5).cuda(), torch.randn(C, generator=g).cuda()
    img_a = torch.zeros(2 * 2 * n * ld, device="cuda")
    img_b = torch.zeros_like(img_a)
    assert dg.lin_bn_relu_jvp_to_image(z, dz, w, C, HW, n // groups, pb, bw, bb, 1e-5, True, (img_a, n))
    torch.backends.cuda.matmul.allow_tf32 = False
    x, t = (z @ w).contiguous(), (dz @ w).contiguous()
    assert dg.bn_relu_jvp(x, t, C, HW, n // groups, pb, bw, bb, 1e-5, True, image=(img_b, n))
    torch.cuda.synchronize()
    a, b = img_a.cpu().numpy(), img_b.cpu().numpy()
    # hi + lo of an element sit 2048 floats apart inside a (128-row, 16-column) tile pair; compare the recombined values
    va = a.reshape(-1, 2, 2048).sum(axis=1)
    vb = b.reshape(-1, 2, 2048).sum(axis=1)
    scale = np.abs(vb).max()
    assert np.abs(va - vb).max() <= 2e-6 * scale
    assert (va != 0).mean() > 0.3  # (ReLU zeroes roughly half)


@pytest.mark.parametrize("rows,K", [(128, 128), (1000, 1024), (65536, 1024)])
def test_last_layer_sigmoid_jvp(rows, K):
    """geocb_lin_sigmoid_jvp against float64: s (1 - s) (t @ w), s = sigmoid(x @ w + b)."""
    from vqvae_b200 import device_graph as dg
    g = torch.Generator().manual_seed(rows + K)
    x = torch.relu(torch.randn(rows, K, generator=g)).cuda()
    t = (torch.randn(rows, K, generator=g) * 0.1).cuda()
    w = (torch.randn(K, 16, generator=g) * 0.05).cuda()
    b = torch.randn(16, generator=g).cuda()
    out = dg.lin_sigmoid_jvp(x, t, w, b)
    assert out is not None and out.shape == (rows, 16)
    y64 = x.double() @ w.double() + b.double()
    s64 = torch.sigmoid(y64)
    ref = s64 * (1 - s64) * (t.double() @ w.double())
    scale = (t.double().abs() @ w.double().abs()).clamp_min(1e-30) * 0.25
    err = ((out.double() - ref).abs() / scale).max().item()
    lib = torch.sigmoid(x @ w + b)
    lib = (lib * (1 - lib) * (t @ w)).double()
    err_lib = ((lib - ref).abs() / scale).max().item()
    print(f"rows={rows} K={K}: fused {err:.3g}, library ops {err_lib:.3g} (of 0.25 sum|t||w|)")
    assert err < 2e-6
    assert dg.lin_sigmoid_jvp(x, t, w[:, :8].contiguous(), None) is None  # unsupported width -> library path
