"""The C-ABI library loads on a machine without a GPU, exports every symbol that
include/b200ft.h declares, the ctypes table covers exactly those, and argument validation
(which needs no device) reports errors instead of crashing."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200ft.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200ft_[a-z0-9_]+)\s*\(", src)))


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    from b200ft import _lib
    return _lib


def test_header_and_ctypes_table_agree(lib):
    assert declared_symbols() == sorted(lib.SIGNATURES)


def test_every_declared_symbol_is_exported(lib):
    handle = ctypes.CDLL(lib.LIB_PATH)
    for name in declared_symbols():
        assert getattr(handle, name) is not None, name
    assert b"sm_100a" in lib.lib().b200ft_version()


def test_gemm_args_struct_layout_matches_header(lib):
    # field order/types of the ctypes mirror vs the header's struct body
    src = open(HEADER).read()
    body = re.search(r"typedef struct b200ft_gemm_args \{(.*?)\} b200ft_gemm_args;", src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    names = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        for part in decl.split(","):
            names.append(re.findall(r"([a-z_0-9]+)\s*$", part.strip())[0])
    assert names == [f[0] for f in lib.GemmArgs._fields_]


def test_argument_validation_needs_no_gpu(lib):
    L = lib.lib()
    assert L.b200ft_adamw_step(None, None, None, None, 0, None, 10, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, 1.0, None, None) == -1
    assert b"adamw_step" in L.b200ft_last_error_string()
    assert L.b200ft_gemm_bf16(None, None) == -1
    args = lib.GemmArgs()
    args.m, args.n, args.k1 = 128, 100, 64          # n not a multiple of 8
    args.a1 = args.b1 = args.c = 4096
    args.lda1 = args.ldb1 = 64
    args.ldc = 104
    assert L.b200ft_gemm_bf16(ctypes.byref(args), None) == -1
    assert b"multiples of 8" in L.b200ft_last_error_string()
    assert L.b200ft_softmax_xent_workspace_bytes(100) >= 400
    assert L.b200ft_softmax_xent_fwd_bwd(4096, 4, 10, 9, 4096, -100, 1.0, 4096, 4096, 1 << 20, None) == -1  # ld < vocab
    assert L.b200ft_layernorm_fwd(4096, 4096, 4096, 4096, 4096, 4096, 4, 12, 1e-5, None) == -1   # d % 8 != 0
    assert L.b200ft_lora_grad(4096, 16, 17, 4096, 64, 64, 8, 4096, 64, 1.0, None) == -1           # r > 16


def test_ops_refuse_cpu_tensors(lib):
    import torch
    from b200ft import ops
    t = torch.zeros(8, 8, dtype=torch.bfloat16)
    with pytest.raises(lib.B200FTError):
        ops.gemm(t, t, t)


def test_engine_refuses_to_run_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from b200ft import gpt2
    with pytest.raises(RuntimeError):
        gpt2.GPT2LoRAStep(gpt2.GPT2Dims(), gpt2.LoRASpec(), gpt2.OptSpec(), {}, {}, batch=1, seq=8)
