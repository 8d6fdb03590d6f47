"""The repo's own causal flash-attention (csrc/attention.cu, csrc/attention_bwd.cu) as a step-engine
mixin.  Engines take ``attention="own"`` (tcgen05 kernels of this library: d_head 64 — GPT-2, Llama-3.2-1B — and
128 — Llama-3-8B: no library kernel in the step), ``attention="sdpa"`` (torch SDPA -> cuDNN, a library: faster today, kept
for A/B measurements and for other head sizes) or ``attention="auto"`` (the default: own wherever it is built).
SURVEY.md §8 a2.8 / (f)-1."""
from __future__ import annotations

import math

import torch

from . import ops


OWN_HEAD_SIZES = (64, 128)


def resolve_attention(mode: str, d_head: int) -> str:
    """"auto" -> "own" for the head sizes the kernels are built for, "sdpa" otherwise; explicit choices pass through."""
    if mode not in ("auto", "own", "sdpa"):
        raise ValueError(f"attention must be auto, own or sdpa, not {mode!r}")
    if mode == "auto":
        return "own" if d_head in OWN_HEAD_SIZES else "sdpa"
    return mode


class OwnAttentionMixin:
    """Expects B, T, M, dev and per-layer fused projection outputs ``self.qkv[l]`` laid out q | k | v."""

    def _own_attn_setup(self, n_head: int, n_kv_head: int, d_head: int, n_layer: int) -> None:
        if d_head not in (64, 128):
            raise ValueError(f"attention='own' is built for d_head 64 and 128 (this model has {d_head}); use 'sdpa'")
        bf, f32 = dict(dtype=torch.bfloat16, device=self.dev), dict(dtype=torch.float32, device=self.dev)
        nq, nkv = n_head * d_head, n_kv_head * d_head
        self._att = dict(batch=self.B, seq=self.T, n_head=n_head, n_kv_head=n_kv_head, d_head=d_head, q_off=0,
                         k_off=nq, v_off=nq + nkv, scale=1.0 / math.sqrt(d_head))
        self._att_cols = (nq, nkv)
        self.att_out = [torch.empty(self.M, nq, **bf) for _ in range(n_layer)]
        self.att_lse = [torch.empty(self.B * n_head, self.T, **f32) for _ in range(n_layer)]
        self.att_ws = ops.attn_bwd_workspace(self.B, self.T, n_head, d_head, self.dev)
        if not hasattr(self, "dqkv"):
            self.dqkv = torch.empty(self.M, nq + 2 * nkv, **bf)

    def _own_attn_fwd(self, l: int) -> torch.Tensor:
        return ops.attn_fwd(self.qkv[l], self.att_out[l], self.att_lse[l], **self._att)

    def _own_attn_bwd(self, l: int, da: torch.Tensor):
        """dq | dk | dv land in self.dqkv; returns the three column-block views."""
        ops.attn_bwd(self.qkv[l], self.att_out[l], da, self.att_lse[l], self.att_ws, self.dqkv, **self._att)
        nq, nkv = self._att_cols
        return self.dqkv[:, :nq], self.dqkv[:, nq: nq + nkv], self.dqkv[:, nq + nkv:]
