"""geocb_pairwise_sum_f32_static -- the allocation-free entry point (scratch in a file-scope `static __device__` array,
csrc/kpp.cu) -- against the existing geocb_pairwise_sum_f32, through the C ABI.  No counterpart in the reference;
the arithmetic it must reproduce is NumPy's float32
pairwise sum (src/geo/kmeans_optimized.py:54,129-134), so the bar is bit equality, as in test_pairwise_sum_matches_numpy.
Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(lib, x, square):
    """(static-scratch result, existing entry point's result) as float32 bit patterns."""
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    a, b, used = C.c_float(-1.0), C.c_float(-2.0), C.c_int64(0)
    rc = lib.geocb_pairwise_sum_f32_static(C.c_void_p(x.data_ptr()), x.numel(), int(square), C.byref(a), stream)
    assert rc == 0, lib.geocb_last_error().decode()
    rc = lib.geocb_pairwise_sum_f32(C.c_void_p(x.data_ptr()), x.numel(), int(square), 0, None, 0, C.byref(b),
                                    C.byref(used), stream)
    assert rc == 0, lib.geocb_last_error().decode()
    return np.float32(a.value), np.float32(b.value)


def test_static_scratch_sum_two_calls_in_a_row():
    assert torch.cuda.is_available()
    from vqvae_b200 import _lib
    lib = _lib.load()
    max_n = lib.geocb_pairwise_sum_f32_static_max_n()
    assert max_n == 4 << 20
    g = torch.Generator().manual_seed(7)
    # first call: the largest size the array serves; second call: a smaller, odd size on the array the first call
    # left dirty (the scratch keeps the last call's contents), squared like the k-means++ weights / the QE
    x1 = (torch.rand(max_n, generator=g) * 3).cuda()
    x2 = torch.randn(1_000_003, generator=g).cuda()
    before = torch.cuda.memory_allocated()
    s1, r1 = _both(lib, x1, square=False)
    s2, r2 = _both(lib, x2, square=True)
    assert torch.cuda.memory_allocated() == before
    assert s1.tobytes() == r1.tobytes() and s2.tobytes() == r2.tobytes(), (s1, r1, s2, r2)
    assert s1 == np.sum(x1.cpu().numpy()) and s2 == np.sum(np.square(x2.cpu().numpy()))
    # a third, tiny call and the empty input
    x3 = torch.arange(5, dtype=torch.float32).cuda()
    s3, r3 = _both(lib, x3, square=False)
    assert s3 == r3 == np.float32(10.0)
    out = C.c_float(-1.0)
    assert lib.geocb_pairwise_sum_f32_static(C.c_void_p(x3.data_ptr()), 0, 0, C.byref(out), None) == 0 and out.value == 0.0


def test_static_scratch_sum_rejects_oversize():
    from vqvae_b200 import _lib
    lib = _lib.load()
    max_n = lib.geocb_pairwise_sum_f32_static_max_n()
    x = torch.zeros(max_n + 1, device="cuda")
    out = C.c_float(-1.0)
    rc = lib.geocb_pairwise_sum_f32_static(C.c_void_p(x.data_ptr()), x.numel(), 0, C.byref(out), None)
    assert rc == -3 and out.value == -1.0  # GEOCB_ERR_WORKSPACE, output untouched
    msg = lib.geocb_last_error().decode()
    assert "exceeds the static scratch limit" in msg and str(max_n) in msg
    # and the array still works afterwards
    s, r = _both(lib, x[:1000] + 0.5, square=True)
    assert s == r == np.float32(250.0)
