"""Llama-family LoRA fine-tuning step on one B200 (BASELINE.json configs[2..4]), written against
the C ABI like gpt2.py.  RMSNorm, rotary embedding, grouped-query attention, SwiGLU, no
biases; LoRA on q_proj and v_proj.  Every GEMM is the tcgen05 kernel: the fused q|k|v
projection carries both adapters as ONE extra K segment (U = s*h*[Aq;Av]^T, 32 wide, against a
block-structured [Nqkv, 32] matrix holding Bq^T and Bv^T), and its backward is a single
four-segment launch (dq.Wq + dk.Wk + dv.Wv + dU.[Aq;Av]).

nn.Linear weights are [out, in], which IS the K-major B operand of forward; the transposed
copies made at load time serve backward.  Attention: torch SDPA (library) or this library's own kernels, as in gpt2.py.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn.functional as F

from . import ops
from ._lib import EPI_NONE, EPI_RESIDUAL
from .attention import OwnAttentionMixin, resolve_attention
from .gpt2 import LoRASpec, OptSpec, R_PAD, StepHostMixin

A_BLK = 2 * R_PAD  # adapter rows per layer in the flat buffers: q adapter 0..15, v adapter 16..31


@dataclass(frozen=True)
class LlamaDims:
    n_layer: int = 16
    d_model: int = 2048
    n_head: int = 32
    n_kv_head: int = 8
    d_ff: int = 8192
    vocab: int = 128256
    ctx: int = 131072
    rms_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_llama3: bool = True
    rope_factor: float = 32.0
    rope_low_freq_factor: float = 1.0
    rope_high_freq_factor: float = 4.0
    rope_original_ctx: int = 8192
    tie_embeddings: bool = True
    d_head_override: int = 0

    @property
    def d_head(self) -> int:
        return self.d_head_override or self.d_model // self.n_head

    @property
    def vocab_padded(self) -> int:
        return (self.vocab + 127) // 128 * 128


LLAMA_3_2_1B = LlamaDims()
LLAMA_3_8B = LlamaDims(n_layer=32, d_model=4096, n_head=32, n_kv_head=8, d_ff=14336, ctx=8192,
                       rope_llama3=False, tie_embeddings=False)


def rope_tables(dims: LlamaDims, seq: int, device) -> tuple[torch.Tensor, torch.Tensor]:
    """cos, sin [seq, d_head/2] fp32; inverse frequencies theta^(-2i/dh) with the Llama-3.1
    wavelength-dependent rescaling ("llama3" rope type) when enabled."""
    dh = dims.d_head
    inv = 1.0 / (dims.rope_theta ** (torch.arange(0, dh, 2, dtype=torch.float64) / dh))
    if dims.rope_llama3:
        wl = 2 * math.pi / inv
        lo, hi = dims.rope_original_ctx / dims.rope_low_freq_factor, dims.rope_original_ctx / dims.rope_high_freq_factor
        smooth = (dims.rope_original_ctx / wl - dims.rope_low_freq_factor) / (
            dims.rope_high_freq_factor - dims.rope_low_freq_factor)
        mid = (1 - smooth) * inv / dims.rope_factor + smooth * inv
        inv = torch.where(wl > lo, inv / dims.rope_factor, torch.where(wl >= hi, mid, inv))
    ang = torch.arange(seq, dtype=torch.float32)[:, None] * inv.float()[None, :]
    return ang.cos().contiguous().to(device), ang.sin().contiguous().to(device)


class LlamaLoRAStep(StepHostMixin, OwnAttentionMixin):
    """Device state of one rank; `step()` = fwd + loss + bwd + AdamW on the q/v adapters."""

    def __init__(self, dims: LlamaDims, lora: LoRASpec, opt: OptSpec, base: dict, adapters: dict, *,
                 batch: int, seq: int, device="cuda:0", world_size: int = 1, grad_sync=None, attention: str = "auto"):
        if not torch.cuda.is_available():
            raise RuntimeError("LlamaLoRAStep needs a CUDA device: there is no CPU path")
        self.attention = resolve_attention(attention, dims.d_head)
        assert lora.r <= R_PAD and dims.d_head % 16 == 0 and dims.d_model % 8 == 0 and dims.d_ff % 8 == 0
        self.dims, self.lora, self.opt = dims, lora, opt
        self.B, self.T, self.M = batch, seq, batch * seq
        self.dev = torch.device(device)
        self.world, self.grad_sync, self.step_count = world_size, grad_sync, 0
        self.Nq, self.Nkv = dims.n_head * dims.d_head, dims.n_kv_head * dims.d_head
        self.Nqkv = self.Nq + 2 * self.Nkv
        self._upload_base(base)
        self._init_adapters(adapters)
        self._alloc()

    # ------------------------------------------------------------------ parameters
    def _bf16(self, t):
        src = t.to(self.dev, torch.float32).contiguous()
        out = torch.empty(src.shape, dtype=torch.bfloat16, device=self.dev)
        ops.cast_f32_to_bf16(src, out)
        return out

    def _f32(self, t):
        return t.to(self.dev, torch.float32).contiguous()

    def _T(self, w):
        out = torch.empty(w.shape[1], w.shape[0], dtype=torch.bfloat16, device=self.dev)
        ops.transpose(w, out)
        return out

    def _upload_base(self, p: dict) -> None:
        D, dev = self.dims, self.dev
        d, Vp = D.d_model, D.vocab_padded
        emb = torch.zeros(Vp, d, dtype=torch.bfloat16, device=dev)
        emb[: D.vocab] = self._bf16(p["embed"])
        self.embed = emb
        if D.tie_embeddings:
            self.head = emb
        else:
            self.head = torch.zeros(Vp, d, dtype=torch.bfloat16, device=dev)
            self.head[: D.vocab] = self._bf16(p["lm_head.w"])
        self.headT = self._T(self.head)                                   # [d, Vp]
        self.layers = []
        for l in range(D.n_layer):
            L = {}
            wqkv = torch.cat([self._bf16(p[f"l{l}.{n}.w"]) for n in ("q_proj", "k_proj", "v_proj")], dim=0)
            L["qkv"], L["qkvT"] = wqkv, self._T(wqkv)                     # [Nqkv, d], [d, Nqkv]
            wo = self._bf16(p[f"l{l}.o_proj.w"])
            L["o"], L["oT"] = wo, self._T(wo)                             # [d, Nq], [Nq, d]
            wgu = torch.cat([self._bf16(p[f"l{l}.gate_proj.w"]), self._bf16(p[f"l{l}.up_proj.w"])], dim=0)
            L["gu"], L["guT"] = wgu, self._T(wgu)                         # [2ff, d], [d, 2ff]
            wd = self._bf16(p[f"l{l}.down_proj.w"])
            L["down"], L["downT"] = wd, self._T(wd)                       # [d, ff], [ff, d]
            L["attn_norm"], L["mlp_norm"] = self._f32(p[f"l{l}.attn_norm.g"]), self._f32(p[f"l{l}.mlp_norm.g"])
            self.layers.append(L)
        self.norm_g = self._f32(p["norm.g"])
        self.cos, self.sin = rope_tables(D, self.T, dev)

    def _init_adapters(self, ad: dict) -> None:
        """Flat fp32 master [At_all (L*32, d) | Bq_all (L*32, Nq) | Bv_all (L*32, Nkv)]; in every
        32-row layer block rows 0..15 belong to the q adapter and 16..31 to the v adapter (rank
        padded to 16, unused rows stay exactly zero).  At = lora_A ([r, in] as PEFT stores it),
        B* = lora_B^T."""
        D, dev = self.dims, self.dev
        L, d, r = D.n_layer, D.d_model, self.lora.r
        nA, nQ, nV = L * A_BLK * d, L * A_BLK * self.Nq, L * A_BLK * self.Nkv
        self.P = torch.zeros(nA + nQ + nV, dtype=torch.float32, device=dev)

        def carve(buf):
            return (buf[:nA].view(L * A_BLK, d), buf[nA:nA + nQ].view(L * A_BLK, self.Nq),
                    buf[nA + nQ:].view(L * A_BLK, self.Nkv))

        self.At_all, self.Bq_all, self.Bv_all = carve(self.P)
        for l in range(L):
            o = l * A_BLK
            self.At_all[o: o + r] = self._f32(ad[f"l{l}.q_proj.lora_A"])
            self.At_all[o + R_PAD: o + R_PAD + r] = self._f32(ad[f"l{l}.v_proj.lora_A"])
            self.Bq_all[o: o + r] = self._f32(ad[f"l{l}.q_proj.lora_B"]).t()
            self.Bv_all[o + R_PAD: o + R_PAD + r] = self._f32(ad[f"l{l}.v_proj.lora_B"]).t()
        self.Mo, self.Vo, self.G = torch.zeros_like(self.P), torch.zeros_like(self.P), torch.zeros_like(self.P)
        self.gAt_all, self.gBq_all, self.gBv_all = carve(self.G)
        self.S = torch.empty(self.P.numel(), dtype=torch.bfloat16, device=dev)
        self.At_sh, self.Bq_sh, self.Bv_sh = carve(self.S)
        self.A_T = torch.empty(d, L * A_BLK, dtype=torch.bfloat16, device=dev)           # [d, L*32]
        self.Bt_blk = torch.zeros(self.Nqkv, L * A_BLK, dtype=torch.bfloat16, device=dev)  # k rows stay zero
        self.norm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.clip = torch.ones(1, dtype=torch.float32, device=dev)
        self.norm_ws = ops.gradnorm_workspace(self.P.numel(), dev)
        self._refresh_shadows(cast=True)

    def _refresh_shadows(self, cast: bool) -> None:
        if cast:
            ops.cast_f32_to_bf16(self.P, self.S)
        ops.transpose(self.At_sh, self.A_T)
        ops.transpose(self.Bq_sh, self.Bt_blk[: self.Nq])
        ops.transpose(self.Bv_sh, self.Bt_blk[self.Nq + self.Nkv:])

    def _export(self, At, Bq, Bv) -> dict:
        r, out = self.lora.r, {}
        for l in range(self.dims.n_layer):
            o = l * A_BLK
            out[f"l{l}.q_proj.lora_A"] = At[o: o + r].contiguous().cpu()
            out[f"l{l}.v_proj.lora_A"] = At[o + R_PAD: o + R_PAD + r].contiguous().cpu()
            out[f"l{l}.q_proj.lora_B"] = Bq[o: o + r].t().contiguous().cpu()
            out[f"l{l}.v_proj.lora_B"] = Bv[o + R_PAD: o + R_PAD + r].t().contiguous().cpu()
        return out

    def adapters_state(self) -> dict:
        return self._export(self.At_all, self.Bq_all, self.Bv_all)

    def grads_state(self) -> dict:
        return self._export(self.gAt_all, self.gBq_all, self.gBv_all)

    def grad_views(self, lo: int, hi: int) -> list:
        """Gradient slices of layers [lo, hi); the whole model is the one flat buffer (a single all-reduce)."""
        if lo == 0 and hi == self.dims.n_layer:
            return [self.G]
        return [t[lo * A_BLK: hi * A_BLK] for t in (self.gAt_all, self.gBq_all, self.gBv_all)]

    # ------------------------------------------------------------------ activations
    def _alloc(self) -> None:
        D, dev, M = self.dims, self.dev, self.M
        d, ff, L, Vp = D.d_model, D.d_ff, D.n_layer, D.vocab_padded
        bf, f32 = dict(dtype=torch.bfloat16, device=dev), dict(dtype=torch.float32, device=dev)
        self.x = [torch.empty(M, d, **bf) for _ in range(L + 1)]
        self.x1 = [torch.empty(M, d, **bf) for _ in range(L)]
        self.h1 = [torch.empty(M, d, **bf) for _ in range(L)]
        self.U = [torch.empty(M, A_BLK, **bf) for _ in range(L)]
        self.qkv = [torch.empty(M, self.Nqkv, **bf) for _ in range(L)]
        self.gu = [torch.empty(M, 2 * ff, **bf) for _ in range(L)]
        self.rs1 = [torch.empty(M, **f32) for _ in range(L)]
        self.rs2 = [torch.empty(M, **f32) for _ in range(L)]
        self.rsf = torch.empty(M, **f32)
        self.h2, self.act = torch.empty(M, d, **bf), torch.empty(M, ff, **bf)
        self.hf = torch.empty(M, d, **bf)
        self.logits = torch.empty(M, Vp, **bf)
        self.targets = torch.empty(self.B, self.T, dtype=torch.int64, device=dev)
        self.loss = torch.zeros(1, **f32)
        self.xent_ws = ops.softmax_xent_workspace(M, dev)
        self.dxs = [torch.empty(M, d, **bf) for _ in range(3)]
        self.dact, self.dgu = torch.empty(M, ff, **bf), torch.empty(M, 2 * ff, **bf)
        self.dln, self.da = torch.empty(M, d, **bf), torch.empty(M, self.Nq, **bf)
        self.dU = torch.empty(M, A_BLK, **bf)
        self.attn_saved = [None] * L
        if self.attention == "own":
            self._own_attn_setup(D.n_head, D.n_kv_head, D.d_head, L)

    # ------------------------------------------------------------------ attention: torch SDPA (library) or own kernels
    def _attn_fwd(self, l: int) -> torch.Tensor:
        if self.attention == "own":
            return self._own_attn_fwd(l)
        B, T, D = self.B, self.T, self.dims
        H, KV, dh = D.n_head, D.n_kv_head, D.d_head
        buf = self.qkv[l]
        q = buf[:, : self.Nq].view(B, T, H, dh).permute(0, 2, 1, 3).detach().requires_grad_(True)
        k = buf[:, self.Nq: self.Nq + self.Nkv].view(B, T, KV, dh).permute(0, 2, 1, 3).detach().requires_grad_(True)
        v = buf[:, self.Nq + self.Nkv:].view(B, T, KV, dh).permute(0, 2, 1, 3).detach().requires_grad_(True)
        with torch.enable_grad():
            o = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=(H != KV))
        self.attn_saved[l] = (q, k, v, o)
        a = o.detach().permute(0, 2, 1, 3)
        if not a.is_contiguous():
            a = a.contiguous()
        return a.view(self.M, self.Nq)

    def _attn_bwd(self, l: int, da: torch.Tensor):
        if self.attention == "own":
            return self._own_attn_bwd(l, da)
        B, T, D = self.B, self.T, self.dims
        q, k, v, o = self.attn_saved[l]
        go = da.view(B, T, D.n_head, D.d_head).permute(0, 2, 1, 3)
        grads = torch.autograd.grad(o, (q, k, v), go)
        self.attn_saved[l] = None
        out = []
        for g, n in zip(grads, (self.Nq, self.Nkv, self.Nkv)):
            g = g.permute(0, 2, 1, 3)
            if not g.is_contiguous():
                g = g.contiguous()
            out.append(g.view(self.M, n))
        return out

    # ------------------------------------------------------------------ the step
    @torch.no_grad()
    def forward_loss(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        D, s = self.dims, self.lora.scale
        ops.embed_fwd(ids.view(-1), self.embed, None, self.x[0], seq=self.T, vocab=D.vocab)
        for l, W in enumerate(self.layers):
            blk = slice(l * A_BLK, (l + 1) * A_BLK)
            ops.rmsnorm_fwd(self.x[l], W["attn_norm"], self.h1[l], self.rs1[l], D.rms_eps)
            ops.gemm(self.h1[l], self.At_sh[blk], self.U[l], alpha=s)
            ops.gemm(self.h1[l], W["qkv"], self.qkv[l], a2=self.U[l], b2=self.Bt_blk[:, blk])
            ops.rope_inplace(self.qkv[l], self.cos, self.sin, seq=self.T, n_heads=D.n_head + D.n_kv_head,
                             d_head=D.d_head)
            a = self._attn_fwd(l)
            ops.gemm(a, W["o"], self.x1[l], epilogue=EPI_RESIDUAL, aux=self.x[l])
            ops.rmsnorm_fwd(self.x1[l], W["mlp_norm"], self.h2, self.rs2[l], D.rms_eps)
            ops.gemm(self.h2, W["gu"], self.gu[l])
            ops.swiglu_fwd(self.gu[l], self.act)
            ops.gemm(self.act, W["down"], self.x[l + 1], epilogue=EPI_RESIDUAL, aux=self.x1[l])
        ops.rmsnorm_fwd(self.x[D.n_layer], self.norm_g, self.hf, self.rsf, D.rms_eps)
        ops.gemm(self.hf, self.head, self.logits)
        ops.shift_targets(labels, self.targets)
        ops.softmax_xent_fwd_bwd(self.logits, self.targets.view(-1), self.loss, self.xent_ws, vocab=D.vocab)
        return self.loss

    @torch.no_grad()
    def backward(self) -> None:
        D, s = self.dims, self.lora.scale
        Nq, Nkv, d = self.Nq, self.Nkv, D.d_model
        ops.zero(self.G)
        dx, dx1, dx0 = self.dxs
        ops.gemm(self.logits, self.headT, self.dln)
        ops.rmsnorm_bwd_dx(self.dln, self.x[D.n_layer], self.norm_g, self.rsf, dx)
        for l in range(D.n_layer - 1, -1, -1):
            W = self.layers[l]
            o = l * A_BLK
            blk = slice(o, o + A_BLK)
            # MLP branch
            ops.gemm(dx, W["downT"], self.dact)
            ops.swiglu_bwd(self.dact, self.gu[l], self.dgu)
            ops.gemm(self.dgu, W["guT"], self.dln)
            ops.rmsnorm_bwd_dx(self.dln, self.x1[l], W["mlp_norm"], self.rs2[l], dx1, dres=dx)
            # attention branch
            ops.gemm(dx1, W["oT"], self.da)
            dq, dk, dv = self._attn_bwd(l, self.da)
            if self.attention == "own":   # dq | dk | dv are one buffer: one launch rotates the q and k heads back
                ops.rope_inplace(self.dqkv, self.cos, self.sin, seq=self.T, n_heads=D.n_head + D.n_kv_head,
                                 d_head=D.d_head, backward=True)
            else:
                ops.rope_inplace(dq, self.cos, self.sin, seq=self.T, n_heads=D.n_head, d_head=D.d_head, backward=True)
                ops.rope_inplace(dk, self.cos, self.sin, seq=self.T, n_heads=D.n_kv_head, d_head=D.d_head, backward=True)
            # adapters: dBq = Uq^T dq, dBv = Uv^T dv, dU = s [dq Bq^T | dv Bv^T], dA = dU^T h1.  The q and v adapters
            # of a layer are stacked (rows 0..15 | 16..31 of the layer's block, the other adapter's rows zero in Bq / Bv),
            # so dU and dA take one launch each for both adapters.
            Uq, Uv = self.U[l][:, :R_PAD], self.U[l][:, R_PAD:]
            ops.lora_grad(Uq, dq, self.gBq_all[o: o + R_PAD], r=R_PAD, cols=Nq)
            ops.lora_grad(Uv, dv, self.gBv_all[o + R_PAD: o + A_BLK], r=R_PAD, cols=Nkv)
            ops.gemm(dq, self.Bq_sh[blk], self.dU, alpha=s, a2=dv, b2=self.Bv_sh[blk])
            ops.lora_grad(self.dU, self.h1[l], self.gAt_all[blk], r=A_BLK, cols=d)
            WT = W["qkvT"]
            ops.gemm(dq, WT[:, :Nq], self.dln, a2=dk, b2=WT[:, Nq: Nq + Nkv],
                     more=[(dv, WT[:, Nq + Nkv:]), (self.dU, self.A_T[:, blk])])
            ops.rmsnorm_bwd_dx(self.dln, self.x[l], W["attn_norm"], self.rs1[l], dx0, dres=dx1)
            dx, dx0 = dx0, dx
            if self.grad_sync is not None:
                self.grad_sync(self, l)

    @torch.no_grad()
    def optimizer_step(self) -> None:
        o = self.opt
        self.step_count += 1
        gs = 1.0 / self.world
        coef = None
        if o.max_grad_norm > 0:
            ops.gradnorm_sq(self.G, self.norm_sq, self.norm_ws)
            ops.clip_coef(self.norm_sq, self.clip, max_norm=o.max_grad_norm, pre_scale=gs)
            coef = self.clip
        ops.adamw_step(self.P, self.Mo, self.Vo, self.G, step=self.step_count, lr=o.lr, beta1=o.beta1,
                       beta2=o.beta2, eps=o.eps, weight_decay=o.weight_decay, shadow=self.S, grad_scale=gs,
                       grad_scale_dev=coef)
        self._refresh_shadows(cast=False)

    def step(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        self.forward_loss(ids, labels)
        self.backward()
        if self.grad_sync is not None:
            self.grad_sync(self, -1)
        self.optimizer_step()
        return self.loss
