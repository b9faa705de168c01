"""ctypes binding of libgeocb.so (C ABI declared in include/geocb.h).

There is no CPU fallback: importing this module works without a GPU (so the
symbol table can be checked), but every compute entry point raises
RuntimeError when the library or a CUDA device is missing.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("GEOCB_LIB", _HERE / "libgeocb.so"))

_p = C.c_void_p
_i64 = C.c_int64
_i32 = C.c_int32
_int = C.c_int
_f64 = C.c_double
_pi64 = C.POINTER(C.c_int64)

# name -> (restype, argtypes); mirrors include/geocb.h one to one
SIGNATURES = {
    "geocb_abi_version": (_int, []),
    "geocb_last_error": (C.c_char_p, []),
    "geocb_device_info": (_int, [C.POINTER(_int), C.POINTER(_int)]),
    "geocb_knn_topk": (_int, [_p, _i64, _int, _i64, _i64, _int, _p, _p, _p, _p]),
    "geocb_knn_query": (_int, [_p, _i64, _p, _i64, _int, _int, _p, _p, _p, _p]),
    "geocb_sym_count": (_int, [_p, _p, _i64, _int, _int, _p, _pi64, _p]),
    "geocb_sym_fill": (_int, [_p, _p, _i64, _int, _int, _p, _p, _p, _p]),
    "geocb_upper_count": (_int, [_p, _p, _i64, _p, _pi64, _p]),
    "geocb_upper_fill": (_int, [_p, _p, _i64, _p, _p, _p, _p, _p, _p]),
    "geocb_edge_len_scatter": (_int, [_p, _p, _i64, _int, _i64, _p, _p, _p, _p, _p]),
    "geocb_bn_relu_jvp": (_int, [_p, _p, _i64, _int, _int, _int, _p, _p, _p, C.c_float, _int, _p]),
    "geocb_bn_relu_jvp_to_image": (_int, [_p, _p, _i64, _int, _int, _int, _p, _p, _p, C.c_float, _int, _p, _i64, _p]),
    "geocb_lin_bn_relu_jvp_to_image": (_int, [_p, _p, _int, _p, _i64, _int, _int, _int, _p, _p, _p, C.c_float, _int, _p, _i64, _p]),
    "geocb_lin_sigmoid_jvp": (_int, [_p, _p, _i64, _int, _int, _p, _p, _p, _p]),
    "geocb_gemm3x_supported": (_int, [_i64, _int, _int]),
    "geocb_gemm3x_prep_weights": (_int, [_p, _int, _int, _p, _p]),
    "geocb_gemm3x": (_int, [_p, _i64, _int, _int, _p, _p, _p, _p]),
    "geocb_cc_labels": (_int, [_p, _p, _i64, _p, _p]),
    "geocb_lcc_mask": (_int, [_p, _i64, _p, _p, _pi64, _pi64, _p]),
    "geocb_csr_compact_count": (_int, [_p, _p, _p, _i64, _p, _pi64, _p]),
    "geocb_csr_compact_fill": (_int, [_p, _p, _p, _p, _p, _i64, _p, _p, _p, _p]),
    "geocb_csr_sym_check": (_int, [_p, _p, _p, _i64, _p, C.POINTER(_int), _p]),
    "geocb_csr_undirected_count": (_int, [_p, _p, _i64, _p, _pi64, _p]),
    "geocb_csr_undirected_fill": (_int, [_p, _p, _p, _i64, _p, _p, _p, _p]),
    "geocb_adj_count": (_int, [_p, _i64, _p, _pi64, _p]),
    "geocb_adj_fill": (_int, [_p, _p, _p, _i64, _int, _p, _p, _p, _p]),
    "geocb_sssp_workspace_bytes": (_i64, [_i64, _int]),
    "geocb_sssp_multi": (_int, [_p, _p, _i64, _p, _int, _f64, _int, _int, _p, _i64, _p, _p, _p, _p, _pi64, _p]),
    "geocb_sssp_vm_slots": (_int, []),
    "geocb_sssp_vm_workspace_bytes": (_i64, [_i64]),
    "geocb_sssp_vm_blocks": (_int, [_p, _p, _i64, _int, _p, _p]),
    "geocb_sssp_multi_vm": (_int, [_p, _p, _int, _i64, _p, _int, _f64, _i64, _p, _i64, _p, _pi64, _p]),
    "geocb_sssp_vm32_workspace_bytes": (_i64, [_i64]),
    "geocb_sssp_multi_vm32": (_int, [_p, _p, _int, _i64, _p, _int, _f64, _i64, _p, _i64, _p, _pi64, _p]),
    "geocb_voronoi_workspace_bytes": (_i64, [_i64]),
    "geocb_voronoi_init": (_int, [_i64, _p, _p, _p, _p, _i64, _p]),
    "geocb_voronoi_insert": (_int, [_p, _p, _i64, _i32, _i32, _f64, _p, _p, _p, _p, _i64, _p]),
    "geocb_voronoi_check": (_int, [_i64, _p, _i64, _p]),
    "geocb_pairwise_plan_size": (_i64, [_i64]),
    "geocb_pairwise_plan_host": (_int, [_i64, _pi64]),
    "geocb_kpp_workspace_bytes": (_i64, [_i64]),
    "geocb_kpp_draw": (_int, [_p, _p, _i64, _f64, _p, _p, _i64, _pi64, _p]),
    "geocb_pairwise_sum_f32": (_int, [_p, _i64, _int, _int, _p, _i64, C.POINTER(C.c_float), _pi64, _p]),
    "geocb_pairwise_sum_f32_static_max_n": (_i64, []),
    "geocb_pairwise_sum_f32_static": (_int, [_p, _i64, _int, C.POINTER(C.c_float), _p]),
}

_lib = None


def load():
    """Load libgeocb.so (once).  Raises RuntimeError if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C vqvae_b200/csrc`; there is no CPU fallback")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    if lib.geocb_abi_version() != 1:
        raise RuntimeError("libgeocb ABI version mismatch")
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().geocb_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")


def require_cuda():
    """The product path needs a CUDA device; never fall back to the CPU."""
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("vqvae_b200 needs a CUDA (sm_100a) device: no CPU fallback exists")
    return torch.device("cuda", torch.cuda.current_device())


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else C.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)
