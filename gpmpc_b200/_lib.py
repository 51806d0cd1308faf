"""ctypes binding of the C ABI declared in include/gpmpc_b200.h.

The shared library is the product: there is no Python / torch / CPU fallback.  If it is missing (not
built) or a call fails, the caller gets an exception -- never a silently different code path.
"""

from __future__ import annotations

import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("GPMPC_B200_LIB") or os.path.join(_HERE, "libgpmpc_b200.so")   # override: A/B builds

_p, _l, _i, _d = C.c_void_p, C.c_long, C.c_int, C.c_double

# name -> (restype, argtypes); must list every symbol include/gpmpc_b200.h declares
SIGNATURES = {
    "gpk_abi_version": (_i, []),
    "gpk_init": (_i, [_p]),
    "gpk_shutdown": (_i, []),
    "gpk_last_error": (C.c_char_p, []),
    "gpk_launch_count": (_l, []),
    "gpk_reset_launch_count": (None, []),
    "gpk_pad_dim": (_i, [_i]),
    "gpk_hyp_stride": (_i, []),
    "gpk_train_tile": (_i, []),
    "gpk_cross_kernel": (_i, [_i, _p, _l, _l, _l, _p, _l, _l, _l, _i, _p, _p, _l, _l, _l, _l, _i, _i, _p]),
    "gpk_potrf": (_i, [_p, _l, _l, _p, _p, _p]),
    "gpk_trtri_scratch_elems": (_l, [_l]),
    "gpk_trtri": (_i, [_p, _l, _l, _p, _p, _l, _p, _l, _p, _l, _p]),
    "gpk_potrf_trtri": (_i, [_p, _l, _l, _p, _p, _p, _l, _p, _l, _p, _l, _p]),
    "gpk_kinv": (_i, [_p, _l, _l, _p, _l, _p]),
    "gpk_kinv_rows": (_i, [_p, _l, _l, _l, _l, _l, _p, _l, _p]),
    "gpk_solve_alpha": (_i, [_p, _l, _p, _l, _p, _l, _l, _p, _p, _p, _p, _p]),
    "gpk_mll_small_max_n": (_i, []),
    "gpk_mll_small_out_stride": (_i, []),
    "gpk_mll_small": (_i, [_p, _l, _l, _p, _l, _l, _i, _p, _p, _p, _l, _p, _p, _l, _l, _i, _p]),
    "gpk_grad_trace_work_elems": (_l, [_l, _i, _l, _l]),
    "gpk_mll_grad_reduce": (_i, [_p, _l, _l, _i, _p, _p, _l, _p, _l, _l, _p, _l, _p, _p]),
    "gpk_mll_grad_reduce_sym": (_i, [_p, _l, _l, _i, _p, _p, _l, _p, _l, _l, _p, _l, _p, _p]),
    "gpk_mll_grad_reduce_static_max_elems": (_l, []),
    "gpk_mll_grad_reduce_static": (_i, [_p, _l, _l, _i, _p, _p, _l, _p, _l, _l, _i, _p, _p]),
    "gpk_pack_train_elems": (_l, [_l, _i]),
    "gpk_pack_train": (_i, [_i, _p, _l, _l, _i, _p, _p, _p, _l, _p]),
    "gpk_mean_jac_work_bytes": (_l, [_i, _l, _l, _i, _i]),
    "gpk_predict_mean_jac": (_i, [_i, _p, _l, _l, _l, _i, _p, _p, _l, _l, _p, _l, _p, _l, _l, _p, _l, _i, _p]),
    "gpk_predict_var_work_bytes": (_l, [_l, _l, _i]),
    "gpk_predict_var": (_i, [_i, _p, _l, _l, _l, _i, _p, _p, _l, _l, _l, _p, _l, _l, _l, _p, _l, _p, _l, _i, _p]),
    "gpk_var_colsq": (_i, [_p, _l, _l, _p, _l, _l, _l, _l, _p, _l, _l, _i, _p]),
    "gpk_var_finalize": (_i, [_i, _p, _l, _l, _i, _l, _p, _p, _l, _i, _p]),
    "gpk_dgemm_nt": (_i, [_p, _l, _p, _l, _p, _l, _l, _l, _l, _d, _d, _i, _i, _p]),
    "gpk_oz_dgemm_nt_work_bytes": (_l, [_l, _l, _l, _i, _i]),
    "gpk_oz_dgemm_nt": (_i, [_p, _l, _p, _l, _p, _l, _l, _l, _l, _d, _d, _i, _i, _i, _p, _l, _p]),
    "gpk_oz_gemm_work_bytes": (_l, [_l, _i]),
    "gpk_oz_gemm_set_workspace": (_i, [_p, _l, _i, _l]),
    "gpk_oz_default_digits": (_i, []),
    "gpk_oz_max_npad": (_l, [_i]),
    "gpk_oz_w_image_bytes": (_l, [_l, _i]),
    "gpk_oz_e_image_bytes": (_l, [_l, _l, _i]),
    "gpk_oz_slice_w": (_i, [_p, _l, _l, _i, _p, _p, _p]),
    "gpk_oz_cross": (_i, [_i, _p, _l, _l, _l, _p, _l, _l, _l, _i, _p, _p, _l, _l, _l, _i, _i, _p]),
    "gpk_predict_i8_work_bytes": (_l, [_l, _l, _i, _i, _i]),
    "gpk_predict_i8": (_i, [_i, _p, _l, _l, _l, _i, _p, _p, _l, _l, _l, _p, _l, _p, _l, _p, _l, _l, _i,
                            _p, _l, _p, _l, _l, _p, _l, _p, _l, _i, _i, _p, _p]),
    "gpk_mean_jac_dmma_work_bytes": (_l, [_l, _i, _i]),
    "gpk_predict_mean_jac_dmma": (_i, [_i, _p, _l, _l, _l, _i, _p, _p, _l, _l, _l, _p, _l, _p, _l, _p, _l, _l, _p, _l, _i, _p]),
    "gpk_oz_var_partials": (_i, [_p, _l, _p, _l, _p, _l, _p, _l, _l, _i, _p, _l, _l, _i, _i, _p, _p]),
    "gpk_oz_query_tile": (_i, []),
    "gpk_oz_pieces": (_i, [_l, _l]),
    "gpk_stage_rk4_work_bytes": (_l, [_l, _l, _i]),
    "gpk_stage_rk4": (_i, [_p, _l, _p, _l, _l, _p, _i, _d, _p, _p, _l, _l, _p, _p, _p, _p, _l, _p]),
    "gpk_tighten_constraints": (_i, [_p, _l, _i, _p, _l, _i, _i, _p, _p, _p, _p, _p, _i, _i, _i, _l, _d, _d, _p, _p, _p]),
}

_lib = None


class GpkError(RuntimeError):
    """A C-ABI call returned a non-zero status."""


def load() -> C.CDLL:
    """Load libgpmpc_b200.so (built by `python __graft_entry__.py` / `make -C gpmpc_b200/csrc`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA extension first (python -c 'import __graft_entry__ as g; "
            "g.build()').  gpmpc_b200 has no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def call(name: str, *args):
    """Call an int-returning entry point and raise GpkError with the library's message on failure."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        raise GpkError(f"{name} failed with status {rc}: {lib.gpk_last_error().decode()}")
    return rc


def ptr(t: torch.Tensor | None) -> int | None:
    if t is None:
        return None
    assert t.is_cuda, "the C ABI takes device pointers"
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(device=None) -> torch.device:
    """The product path needs a CUDA device; fail loudly otherwise (no CPU fallback)."""
    if not torch.cuda.is_available():
        raise RuntimeError("gpmpc_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback.")
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    if dev.type != "cuda":
        raise RuntimeError(f"gpmpc_b200 computes on CUDA only, got device {dev}")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    return dev
