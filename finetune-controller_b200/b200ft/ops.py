"""Thin torch-tensor front ends for the C ABI: they only pull out device pointers, leading
dimensions and the current stream.  torch is plumbing here (memory + streams), not compute.

`LAUNCHES` counts kernels of THIS library enqueued through these wrappers (bench.py reports
it as `gpu_launches`).
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from ._lib import (A_MN_MAJOR, B_MN_MAJOR, EPI_ATOMIC_F32_T, EPI_DGELU, EPI_GELU, EPI_NONE, EPI_RESIDUAL,
                   GemmArgs, check, lib)

LAUNCHES = 0

# Optional per-kernel-class timing (bench.py's instrumented pass): when PROFILE is a dict,
# every wrapper brackets its launch with CUDA events on the launching stream and appends
# (start, end, algorithmic_work) under its class name.  Off (None) in normal operation.
PROFILE = None
NUM_SMS = 148  # B200; mirrors B200FT_NUM_SMS of the library


class _timed:
    def __init__(self, name, work, label=None):
        self.name, self.work, self.label = name, work, label

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e1 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        return self

    def __exit__(self, *exc):
        if PROFILE is not None:
            self.e1.record()
            PROFILE.setdefault(self.name, []).append((self.e0, self.e1, self.work, self.label))
        return False


def _p(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise _lib.B200FTError("b200ft ops take CUDA tensors only (no CPU path exists)")


def gemm(a, b, out, *, a_mn=False, b_mn=False, a2=None, b2=None, more=(), bias=None, epilogue=EPI_NONE,
         aux=None, out2=None, alpha=1.0, block_n=0, max_ctas=0, split_k=0, cta_group=0, m=None, n=None, k=None, k2=None):
    """out[m,n] = epi(alpha * (A.B^T + A2.B2^T + ...) + bias).  A: [M,K] (K-major) or [K,M] (a_mn);
    B: [N,K] or [K,N] (b_mn); A2 [M,K2], B2 [N,K2]; `more` = up to two further (A_i, B_i) pairs.
    Row pitch = stride(0)."""
    global LAUNCHES
    _need_cuda(a, b, out, a2, b2, bias, aux, out2)
    assert a.dtype == b.dtype == torch.bfloat16
    assert out.dtype == (torch.float32 if epilogue == EPI_ATOMIC_F32_T else torch.bfloat16)
    assert a.stride(-1) == 1 and b.stride(-1) == 1 and out.stride(-1) == 1
    if m is None:
        m = a.shape[1] if a_mn else a.shape[0]
    if k is None:
        k = a.shape[0] if a_mn else a.shape[1]
    if n is None:
        n = b.shape[1] if b_mn else b.shape[0]
    args = GemmArgs()
    args.m, args.n = m, n
    args.a1, args.lda1 = a.data_ptr(), a.stride(0)
    args.b1, args.ldb1 = b.data_ptr(), b.stride(0)
    args.k1 = k
    if a2 is not None:
        assert a2.dtype == b2.dtype == torch.bfloat16
        args.a2, args.lda2 = a2.data_ptr(), a2.stride(0)
        args.b2, args.ldb2 = b2.data_ptr(), b2.stride(0)
        args.k2 = k2 if k2 is not None else a2.shape[1]
    kk_more = 0
    for i, (ai, bi) in enumerate(more):
        assert a2 is not None and i < 2 and ai.dtype == bi.dtype == torch.bfloat16
        _need_cuda(ai, bi)
        setattr(args, f"a{i + 3}", ai.data_ptr()); setattr(args, f"lda{i + 3}", ai.stride(0))
        setattr(args, f"b{i + 3}", bi.data_ptr()); setattr(args, f"ldb{i + 3}", bi.stride(0))
        setattr(args, f"k{i + 3}", ai.shape[1])
        kk_more += ai.shape[1]
    args.layout = (A_MN_MAJOR if a_mn else 0) | (B_MN_MAJOR if b_mn else 0)
    args.epilogue = epilogue
    args.alpha = alpha
    if bias is not None:
        assert bias.dtype == torch.float32
        args.bias = bias.data_ptr()
    if aux is not None:
        args.aux, args.ld_aux = aux.data_ptr(), aux.stride(0)
    args.c, args.ldc = out.data_ptr(), out.stride(0)
    if out2 is not None:
        args.c2, args.ldc2 = out2.data_ptr(), out2.stride(0)
    args.block_n, args.max_ctas, args.split_k, args.cta_group = block_n, max_ctas, split_k, cta_group
    kk = k + (args.k2 if a2 is not None else 0) + kk_more
    # which instantiation the library will pick (mirrors b200ft_gemm_bf16's rule for cta_group = 0)
    cg = cta_group or (2 if (epilogue != EPI_ATOMIC_F32_T and m > 128 and n >= 128) else 1)
    with _timed("gemm", 2.0 * m * n * kk, (m, n, kk, int(a_mn), int(b_mn), epilogue, cg)):
        check(lib().b200ft_gemm_bf16(ctypes.byref(args), _stream()), "gemm_bf16")
    LAUNCHES += 1
    return out


def lora_grad(s, t, out, *, r, cols, alpha=1.0, m=None):
    """out[i,c] += alpha * sum_m s[m,i] t[m,c]  (fp32 out, bf16 s/t)."""
    global LAUNCHES
    _need_cuda(s, t, out)
    assert s.dtype == t.dtype == torch.bfloat16 and out.dtype == torch.float32
    m = s.shape[0] if m is None else m
    with _timed("lora_grad", 2.0 * m * (cols + r)):
        check(lib().b200ft_lora_grad(_p(s), s.stride(0), r, _p(t), t.stride(0), cols, m, _p(out), out.stride(0),
                                     alpha, _stream()), "lora_grad")
    LAUNCHES += 1
    return out


def softmax_xent_workspace(rows, device):
    nbytes = lib().b200ft_softmax_xent_workspace_bytes(rows)
    return torch.empty(nbytes, dtype=torch.uint8, device=device)


def softmax_xent_fwd_bwd(logits, targets, loss_out, workspace, *, vocab, grad_scale=1.0, ignore_index=-100):
    """In place: logits [rows, ld] bf16 -> d loss / d logits; loss_out[0] = mean loss."""
    global LAUNCHES
    _need_cuda(logits, targets, loss_out, workspace)
    assert logits.dtype == torch.bfloat16 and targets.dtype == torch.int64 and loss_out.dtype == torch.float32
    rows = logits.shape[0]
    with _timed("softmax_xent", 2.0 * rows * vocab * 2 + 12.0 * rows):
        check(lib().b200ft_softmax_xent_fwd_bwd(_p(logits), rows, vocab, logits.stride(0), _p(targets),
                                                ignore_index, grad_scale, _p(loss_out), _p(workspace),
                                                workspace.numel(), _stream()), "softmax_xent_fwd_bwd")
    LAUNCHES += 2
    return loss_out


def softmax_xent_bad_targets(workspace) -> int:
    """Targets of the LAST softmax_xent_fwd_bwd call on this workspace that were neither ignore_index nor inside
    [0, vocab): the kernel treats them as ignored and counts them in scratch word 2.  Synchronises (reads 4 bytes)."""
    return int(workspace[8:12].view(torch.int32).item())


def shift_targets(labels, targets, ignore_index=-100):
    global LAUNCHES
    _need_cuda(labels, targets)
    b, t = labels.shape
    check(lib().b200ft_shift_targets(_p(labels), _p(targets), b, t, ignore_index, _stream()), "shift_targets")
    LAUNCHES += 1
    return targets


def adamw_step(p, m, v, g, *, step, lr, beta1, beta2, eps, weight_decay, shadow=None, grad_scale=1.0,
               grad_scale_dev=None, max_blocks=0):
    global LAUNCHES
    _need_cuda(p, m, v, g, shadow, grad_scale_dev)
    assert p.dtype == m.dtype == v.dtype == torch.float32
    assert g.dtype in (torch.bfloat16, torch.float32)
    n = p.numel()
    assert m.numel() == n and v.numel() == n and g.numel() == n
    per = 24 + g.element_size() + (2 if shadow is not None else 0)
    with _timed("adamw", float(per) * n):
        check(lib().b200ft_adamw_step_limited(_p(p), _p(m), _p(v), _p(g), int(g.dtype == torch.bfloat16), _p(shadow),
                                              n, lr, beta1, beta2, eps, weight_decay, step, grad_scale,
                                              _p(grad_scale_dev), max_blocks, _stream()), "adamw_step")
    LAUNCHES += 1


def gradnorm_workspace(n, device):
    return torch.zeros(lib().b200ft_gradnorm_sq_workspace_bytes(n), dtype=torch.uint8, device=device)


def gradnorm_sq(g, out_sq, workspace):
    global LAUNCHES
    _need_cuda(g, out_sq, workspace)
    check(lib().b200ft_gradnorm_sq(_p(g), int(g.dtype == torch.bfloat16), g.numel(), _p(out_sq), _p(workspace),
                                   workspace.numel(), _stream()), "gradnorm_sq")
    LAUNCHES += 1
    return out_sq


def clip_coef(norm_sq, coef, *, max_norm, pre_scale=1.0):
    global LAUNCHES
    check(lib().b200ft_clip_coef(_p(norm_sq), max_norm, pre_scale, _p(coef), _stream()), "clip_coef")
    LAUNCHES += 1
    return coef


def layernorm_fwd(x, gamma, beta, y, mean, rstd, eps):
    global LAUNCHES
    _need_cuda(x, gamma, beta, y, mean, rstd)
    rows, d = x.shape
    assert x.is_contiguous() and y.is_contiguous()
    with _timed("layernorm_fwd", 4.0 * rows * d):
        check(lib().b200ft_layernorm_fwd(_p(x), _p(gamma), _p(beta), _p(y), _p(mean), _p(rstd), rows, d, eps,
                                         _stream()), "layernorm_fwd")
    LAUNCHES += 1
    return y


def layernorm_bwd_dx(dy, x, gamma, mean, rstd, dx, dres=None):
    global LAUNCHES
    _need_cuda(dy, x, gamma, mean, rstd, dx, dres)
    rows, d = x.shape
    assert dy.is_contiguous() and x.is_contiguous() and dx.is_contiguous()
    with _timed("layernorm_bwd", (8.0 if dres is not None else 6.0) * rows * d):
        check(lib().b200ft_layernorm_bwd_dx(_p(dy), _p(x), _p(gamma), _p(mean), _p(rstd), _p(dres), _p(dx), rows,
                                            d, _stream()), "layernorm_bwd_dx")
    LAUNCHES += 1
    return dx


def rmsnorm_fwd(x, gamma, y, rstd, eps):
    global LAUNCHES
    _need_cuda(x, gamma, y, rstd)
    rows, d = x.shape
    assert x.is_contiguous() and y.is_contiguous()
    with _timed("rmsnorm_fwd", 4.0 * rows * d):
        check(lib().b200ft_rmsnorm_fwd(_p(x), _p(gamma), _p(y), _p(rstd), rows, d, eps, _stream()), "rmsnorm_fwd")
    LAUNCHES += 1
    return y


def rmsnorm_bwd_dx(dy, x, gamma, rstd, dx, dres=None):
    global LAUNCHES
    _need_cuda(dy, x, gamma, rstd, dx, dres)
    rows, d = x.shape
    assert dy.is_contiguous() and x.is_contiguous() and dx.is_contiguous()
    with _timed("rmsnorm_bwd", (8.0 if dres is not None else 6.0) * rows * d):
        check(lib().b200ft_rmsnorm_bwd_dx(_p(dy), _p(x), _p(gamma), _p(rstd), _p(dres), _p(dx), rows, d, _stream()),
              "rmsnorm_bwd_dx")
    LAUNCHES += 1
    return dx


def rope_inplace(x, cos, sin, *, seq, n_heads, d_head, backward=False, rows=None):
    """Rotate the first n_heads*d_head columns of each row of x (bf16, pitch x.stride(0)) in place."""
    global LAUNCHES
    _need_cuda(x, cos, sin)
    assert x.dtype == torch.bfloat16 and cos.dtype == sin.dtype == torch.float32 and x.stride(-1) == 1
    rows = x.shape[0] if rows is None else rows
    with _timed("rope", 4.0 * rows * n_heads * d_head):
        check(lib().b200ft_rope_inplace(_p(x), x.stride(0), rows, seq, n_heads, d_head, _p(cos), _p(sin),
                                        int(backward), _stream()), "rope_inplace")
    LAUNCHES += 1
    return x


def swiglu_fwd(gate_up, act):
    global LAUNCHES
    _need_cuda(gate_up, act)
    rows, ff = act.shape
    assert gate_up.shape == (rows, 2 * ff) and gate_up.is_contiguous() and act.is_contiguous()
    with _timed("swiglu_fwd", 6.0 * rows * ff):
        check(lib().b200ft_swiglu_fwd(_p(gate_up), _p(act), rows, ff, _stream()), "swiglu_fwd")
    LAUNCHES += 1
    return act


def swiglu_bwd(dact, gate_up, dgate_up, act=None):
    """dgate_up from (dact, gate_up); with `act` given, also rewrites the forward product into it."""
    global LAUNCHES
    _need_cuda(dact, gate_up, dgate_up, act)
    rows, ff = dact.shape
    assert gate_up.shape == (rows, 2 * ff) and dgate_up.shape == (rows, 2 * ff)
    assert dact.is_contiguous() and gate_up.is_contiguous() and dgate_up.is_contiguous()
    assert act is None or (act.shape == (rows, ff) and act.is_contiguous())
    with _timed("swiglu_bwd", (12.0 if act is not None else 10.0) * rows * ff):
        check(lib().b200ft_swiglu_bwd_act(_p(dact), _p(gate_up), _p(dgate_up), _p(act), rows, ff, _stream()),
              "swiglu_bwd")
    LAUNCHES += 1
    return dgate_up


def attn_fwd(qkv, out, lse2, *, batch, seq, n_head, n_kv_head, d_head, q_off, k_off, v_off, scale):
    """Causal attention forward over the fused projection buffer (see include/b200ft.h)."""
    global LAUNCHES
    _need_cuda(qkv, out, lse2)
    assert qkv.dtype == out.dtype == torch.bfloat16 and lse2.dtype == torch.float32
    flops = 4.0 * batch * n_head * d_head * seq * (seq + 128) / 2  # causal, counted in whole 128-key tiles
    with _timed("attn_fwd", flops):
        check(lib().b200ft_attn_fwd(_p(qkv), qkv.stride(0), q_off, k_off, v_off, _p(out), out.stride(0), _p(lse2),
                                    batch, seq, n_head, n_kv_head, d_head, scale, _stream()), "attn_fwd")
    LAUNCHES += 1
    return out


def attn_bwd_workspace(batch, seq, n_head, d_head, device):
    return torch.empty(lib().b200ft_attn_bwd_workspace_bytes(batch, seq, n_head, d_head), dtype=torch.uint8, device=device)


def attn_bwd(qkv, out, dout, lse2, workspace, dqkv, *, batch, seq, n_head, n_kv_head, d_head, q_off, k_off, v_off, scale,
             deterministic=False):
    """Causal attention backward: dq | dk | dv into `dqkv` at the fused projection's column offsets."""
    global LAUNCHES
    _need_cuda(qkv, out, dout, lse2, workspace, dqkv)
    assert qkv.dtype == out.dtype == dout.dtype == dqkv.dtype == torch.bfloat16
    assert lse2.dtype == torch.float32 and workspace.dtype == torch.uint8
    flops = 10.0 * batch * n_head * d_head * seq * (seq + 128) / 2  # five tile GEMMs per visible pair
    with _timed("attn_bwd", flops):
        check(lib().b200ft_attn_bwd(_p(qkv), qkv.stride(0), q_off, k_off, v_off, _p(out), out.stride(0), _p(dout),
                                    dout.stride(0), _p(lse2), _p(workspace), workspace.numel(), _p(dqkv), dqkv.stride(0),
                                    batch, seq, n_head, n_kv_head, d_head, scale,
                                    _lib.ATTN_BWD_DETERMINISTIC if deterministic else 0, _stream()), "attn_bwd")
    LAUNCHES += 3 if (deterministic or d_head == 128) else 2   # delta + (dK|dV, dQ) or delta + fused
    return dqkv


def embed_fwd(ids, wte, wpe, x, *, seq, vocab):
    global LAUNCHES
    _need_cuda(ids, wte, wpe, x)
    rows, d = x.shape
    assert wte.stride(0) == d and (wpe is None or wpe.stride(0) == d)
    check(lib().b200ft_embed_fwd(_p(ids), _p(wte), _p(wpe), _p(x), rows, seq, d, vocab, _stream()), "embed_fwd")
    LAUNCHES += 1
    return x


def rmsnorm_bwd_dgamma(dy, x, rstd, dgamma):
    """dgamma (fp32, accumulated) += column sums of dy * x * rstd."""
    global LAUNCHES
    _need_cuda(dy, x, rstd, dgamma)
    rows, d = x.shape
    assert dy.is_contiguous() and x.is_contiguous() and dgamma.dtype == torch.float32 and dgamma.numel() == d
    with _timed("rmsnorm_dgamma", 4.0 * rows * d):
        check(lib().b200ft_rmsnorm_bwd_dgamma(_p(dy), _p(x), _p(rstd), _p(dgamma), rows, d, _stream()),
              "rmsnorm_bwd_dgamma")
    LAUNCHES += 1


def embed_bwd(ids, dx, dwte, *, vocab):
    """dwte[ids] += dx: rows of one id summed in fp32, added to the bf16 table row once (no atomics, reproducible)."""
    global LAUNCHES
    _need_cuda(ids, dx, dwte)
    rows, d = dx.shape
    assert ids.dtype == torch.int64 and ids.numel() == rows and dx.is_contiguous() and dwte.stride(0) == d
    with _timed("embed_bwd", 6.0 * rows * d):
        check(lib().b200ft_embed_bwd(_p(ids), _p(dx), _p(dwte), rows, d, vocab, _stream()), "embed_bwd")
    LAUNCHES += 1


def transpose(inp, out, rows=None, cols=None):
    """out[c, r] = inp[r, c] (bf16)."""
    global LAUNCHES
    _need_cuda(inp, out)
    rows = inp.shape[0] if rows is None else rows
    cols = inp.shape[1] if cols is None else cols
    check(lib().b200ft_transpose_bf16(_p(inp), inp.stride(0), _p(out), out.stride(0), rows, cols, _stream()),
          "transpose_bf16")
    LAUNCHES += 1
    return out


def zero(t):
    """Stream-ordered zero fill of a contiguous tensor through the C ABI (a DMA memset: no kernel, no torch op)."""
    _need_cuda(t)
    assert t.is_contiguous()
    check(lib().b200ft_zero(_p(t), t.numel() * t.element_size(), _stream()), "zero")
    return t


def set_dependent_launch(enabled: bool) -> bool:
    """Programmatic dependent launch on / off for every later launch of this process; returns the previous setting."""
    return bool(lib().b200ft_set_dependent_launch(int(bool(enabled))))


def cast_f32_to_bf16(inp, out):
    global LAUNCHES
    _need_cuda(inp, out)
    assert inp.dtype == torch.float32 and out.dtype == torch.bfloat16 and inp.numel() == out.numel()
    check(lib().b200ft_cast_f32_to_bf16(_p(inp), _p(out), inp.numel(), _stream()), "cast_f32_to_bf16")
    LAUNCHES += 1
    return out
