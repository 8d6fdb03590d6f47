"""ctypes binding of libb200ft.so (the C ABI declared in include/b200ft.h).

There is no fallback: if the shared library has not been built (``__graft_entry__.build()``, i.e. one
``nvcc -gencode arch=compute_100a,code=sm_100a -shared`` over ``csrc/*.cu``) importing an op raises, and every op raises on
a non-zero return code with the library's own error string.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200ft.so")

EPI_NONE, EPI_RESIDUAL, EPI_GELU, EPI_DGELU, EPI_ATOMIC_F32_T = 0, 1, 2, 3, 4
A_MN_MAJOR, B_MN_MAJOR = 1, 2
ATTN_BWD_DETERMINISTIC = 1


class GemmArgs(ctypes.Structure):
    """Mirror of ``struct b200ft_gemm_args`` (include/b200ft.h)."""
    _fields_ = [
        ("m", c_int32), ("n", c_int32),
        ("a1", c_void_p), ("lda1", c_int64),
        ("b1", c_void_p), ("ldb1", c_int64),
        ("k1", c_int32),
        ("a2", c_void_p), ("lda2", c_int64),
        ("b2", c_void_p), ("ldb2", c_int64),
        ("k2", c_int32),
        ("a3", c_void_p), ("lda3", c_int64),
        ("b3", c_void_p), ("ldb3", c_int64),
        ("k3", c_int32),
        ("a4", c_void_p), ("lda4", c_int64),
        ("b4", c_void_p), ("ldb4", c_int64),
        ("k4", c_int32),
        ("layout", c_int32),
        ("epilogue", c_int32),
        ("alpha", c_float),
        ("bias", c_void_p),
        ("aux", c_void_p), ("ld_aux", c_int64),
        ("c", c_void_p), ("ldc", c_int64),
        ("c2", c_void_p), ("ldc2", c_int64),
        ("block_n", c_int32),
        ("max_ctas", c_int32),
        ("split_k", c_int32),
        ("cta_group", c_int32),
    ]


# name -> (restype, argtypes); every symbol include/b200ft.h declares
SIGNATURES = {
    "b200ft_version": (c_char_p, []),
    "b200ft_last_error_string": (c_char_p, []),
    "b200ft_gemm_bf16": (c_int, [POINTER(GemmArgs), c_void_p]),
    "b200ft_lora_grad": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_int, c_int64, c_void_p,
                                 c_int64, c_float, c_void_p]),
    "b200ft_softmax_xent_workspace_bytes": (c_size_t, [c_int64]),
    "b200ft_softmax_xent_fwd_bwd": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_int64, c_float,
                                            c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200ft_softmax_xent_noalloc_max_rows": (c_int64, []),
    "b200ft_softmax_xent_fwd_bwd_noalloc": (c_int, [c_void_p, c_int64, c_int, c_int64, c_void_p, c_int64, c_float,
                                                    c_void_p, c_void_p]),
    "b200ft_shift_targets": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int64, c_void_p]),
    "b200ft_adamw_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int64, c_float,
                                  c_float, c_float, c_float, c_float, c_int, c_float, c_void_p, c_void_p]),
    "b200ft_adamw_step_limited": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_int64, c_float,
                                          c_float, c_float, c_float, c_float, c_int, c_float, c_void_p, c_int, c_void_p]),
    "b200ft_gradnorm_sq_workspace_bytes": (c_size_t, [c_int64]),
    "b200ft_gradnorm_sq": (c_int, [c_void_p, c_int, c_int64, c_void_p, c_void_p, c_size_t, c_void_p]),
    "b200ft_clip_coef": (c_int, [c_void_p, c_float, c_float, c_void_p, c_void_p]),
    "b200ft_layernorm_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                     c_int, c_float, c_void_p]),
    "b200ft_layernorm_bwd_dx": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                        c_int64, c_int, c_void_p]),
    "b200ft_rmsnorm_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_void_p]),
    "b200ft_rmsnorm_bwd_dx": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int,
                                      c_void_p]),
    "b200ft_rope_inplace": (c_int, [c_void_p, c_int64, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_int,
                                    c_void_p]),
    "b200ft_swiglu_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b200ft_swiglu_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b200ft_swiglu_bwd_act": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b200ft_attn_fwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_int, c_int,
                                c_int, c_int, c_int, c_float, c_void_p]),
    "b200ft_attn_bwd_workspace_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "b200ft_attn_bwd": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_int64, c_void_p, c_int64, c_void_p,
                                c_void_p, c_size_t, c_void_p, c_int64, c_int, c_int, c_int, c_int, c_int, c_float, c_int,
                                c_void_p]),
    "b200ft_embed_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p]),
    "b200ft_rmsnorm_bwd_dgamma": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "b200ft_embed_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "b200ft_transpose_bf16": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "b200ft_cast_f32_to_bf16": (c_int, [c_void_p, c_void_p, c_int64, c_void_p]),
    "b200ft_zero": (c_int, [c_void_p, c_size_t, c_void_p]),
    "b200ft_set_dependent_launch": (c_int, [c_int]),
}

_lib = None


class B200FTError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200FTError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU or PyTorch fallback for the training step)")
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)  # AttributeError if the .so is stale
            fn.restype, fn.argtypes = res, args
        _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().b200ft_last_error_string().decode()
        raise B200FTError(f"{what} failed (rc={rc}): {msg}")
