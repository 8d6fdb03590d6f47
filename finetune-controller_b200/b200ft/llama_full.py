"""Llama-family FULL-parameter fine-tuning step on one B200 (BASELINE.json configs[4]: "Llama-3-8B
full-parameter fine-tune bf16 with fused AdamW"), written against the C ABI like llama.py.

Every matrix is trainable.  All of them live in ONE flat group so that the optimizer is a single
HBM-bound launch (28 B per parameter):

    P  fp32 master      Mo, Vo  fp32 Adam moments      S  bf16 shadow (what the GEMMs read)
    G  bf16 gradients (what the weight-gradient GEMMs write and the all-reduce sums)

laid out  [embed | layer 0: qkv, o, gate|up, down | layer 1 ... | lm_head (untied only)]  so a layer's
gradients are one contiguous slice for the bucketed all-reduce.  The RMSNorm gains form a second,
tiny fp32 group (Pn/Mn/Vn/Gn) updated by a second AdamW launch.

GEMMs per layer: 4 forward, 4 dX (B operand MN-major: the nn.Linear weight [out, in] is read as
[K, N] without a transposed copy), 4 dW (both operands MN-major: dW[out, in] = dY^T . X with the
token index as K).  All are the cta_group::2 tcgen05 kernel.  Activations kept per layer: x, x1,
qkv (post-RoPE), the attention output and gate|up; the normalised inputs h1/h2 and the SwiGLU
product are recomputed in backward (HBM-bound, ~1 % of the step) instead of stored.
Attention: this library's own kernels (d_head 64 / 128) or torch SDPA (library), as in llama.py.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F

from . import ops
from .attention import OwnAttentionMixin, resolve_attention
from ._lib import EPI_RESIDUAL
from .gpt2 import OptSpec, StepHostMixin
from .llama import LlamaDims, rope_tables

_MATS = ("qkv", "o", "gu", "down")


class LlamaFullStep(StepHostMixin, OwnAttentionMixin):
    """Device state of one rank; `step()` = fwd + loss + bwd (dX and dW) + AdamW on every tensor.

    `base` is a dict in the oracle's naming (oracle/llama_lora_ref.py init_base_params) or None, in
    which case the parameters are drawn in place on the device (normal(0, init_std), unit gains):
    an 8B model never exists twice in memory."""

    def __init__(self, dims: LlamaDims, opt: OptSpec, base: dict | None, *, batch: int, seq: int,
                 device="cuda:0", world_size: int = 1, grad_sync=None, seed: int = 0, init_std: float = 0.02,
                 attention: str = "auto", overlap_optimizer: bool = False):
        if not torch.cuda.is_available():
            raise RuntimeError("LlamaFullStep needs a CUDA device: there is no CPU path")
        self.attention = resolve_attention(attention, dims.d_head)
        assert dims.d_head % 16 == 0 and dims.d_model % 8 == 0 and dims.d_ff % 8 == 0
        self.dims, self.opt, self.lora = dims, opt, None
        self.B, self.T, self.M = batch, seq, batch * seq
        self.dev = torch.device(device)
        self.world, self.grad_sync, self.step_count = world_size, grad_sync, 0
        # SMs left to the overlapped all-reduce in backward (ddp.overlap_group): the GEMM works on CTA
        # pairs, so each reserved collective CTA may cost a whole TPC
        reserve = getattr(grad_sync, "reserve_ctas", 0) if grad_sync is not None else 0
        self.bwd_ctas = 0 if reserve <= 0 else max(2, ops.NUM_SMS - 2 * reserve)
        # Option (off: measured slower): the HBM-bound AdamW of layer l on a second stream as soon as that layer's
        # gradients are final (and, with data parallelism, reduced), beside the backward GEMMs of the layers below.
        # Needs the final gradient per layer: no global-norm clipping, one all-reduce bucket per layer.
        # Llama-3-8B, 1 GPU: 247.7 ms/step with it vs 234.3 without — the chip sits at its power cap during the
        # GEMMs, so the co-running update only lowers their clock and pollutes L2 (GEMMs 1174 -> 899 TFLOP/s).
        per_layer_sync = grad_sync is None or len(getattr(grad_sync, "buckets", ())) == dims.n_layer
        self.overlap_opt = bool(overlap_optimizer) and opt.max_grad_norm <= 0 and per_layer_sync
        self.opt_stream = torch.cuda.Stream(device=self.dev) if self.overlap_opt else None
        self.Nq, self.Nkv = dims.n_head * dims.d_head, dims.n_kv_head * dims.d_head
        self.Nqkv = self.Nq + 2 * self.Nkv
        self._layout()
        self._init_params(base, seed, init_std)
        self._alloc()

    # ------------------------------------------------------------------ parameters
    def _layout(self) -> None:
        D = self.dims
        d, ff, Vp = D.d_model, D.d_ff, D.vocab_padded
        shapes = {"qkv": (self.Nqkv, d), "o": (d, self.Nq), "gu": (2 * ff, d), "down": (d, ff)}
        self.slots = {}          # name -> (offset, shape) into the flat matrix group
        off = 0
        self.slots["embed"] = (off, (Vp, d)); off += Vp * d
        self.layer_off = []
        for l in range(D.n_layer):
            self.layer_off.append(off)
            for n in _MATS:
                self.slots[f"l{l}.{n}"] = (off, shapes[n]); off += shapes[n][0] * shapes[n][1]
        self.layer_off.append(off)
        if not D.tie_embeddings:
            self.slots["head"] = (off, (Vp, d)); off += Vp * d
        self.n_mat = off
        self.n_norm = (2 * D.n_layer + 1) * d

    def _view(self, flat: torch.Tensor, name: str) -> torch.Tensor:
        off, shape = self.slots[name]
        return flat[off: off + shape[0] * shape[1]].view(shape)

    def _init_params(self, base: dict | None, seed: int, std: float) -> None:
        D, dev = self.dims, self.dev
        d, ff, L = D.d_model, D.d_ff, D.n_layer
        f32 = dict(dtype=torch.float32, device=dev)
        self.P = torch.zeros(self.n_mat, **f32)
        self.Pn = torch.ones(self.n_norm, **f32)
        if base is None:
            g = torch.Generator(device=dev).manual_seed(seed)
            for name, (off, shape) in self.slots.items():
                rows = D.vocab if name in ("embed", "head") else shape[0]   # padded vocabulary rows stay zero
                self.P[off: off + rows * shape[1]].normal_(0.0, std, generator=g)
        else:
            def put(name, t, r0=0):
                v = self._view(self.P, name)
                v[r0: r0 + t.shape[0]].copy_(t.to(dev, torch.float32))

            put("embed", base["embed"])
            if not D.tie_embeddings:
                put("head", base["lm_head.w"])
            for l in range(L):
                put(f"l{l}.qkv", base[f"l{l}.q_proj.w"], 0)
                put(f"l{l}.qkv", base[f"l{l}.k_proj.w"], self.Nq)
                put(f"l{l}.qkv", base[f"l{l}.v_proj.w"], self.Nq + self.Nkv)
                put(f"l{l}.o", base[f"l{l}.o_proj.w"])
                put(f"l{l}.gu", base[f"l{l}.gate_proj.w"], 0)
                put(f"l{l}.gu", base[f"l{l}.up_proj.w"], ff)
                put(f"l{l}.down", base[f"l{l}.down_proj.w"])
                self.Pn[2 * l * d: (2 * l + 1) * d].copy_(base[f"l{l}.attn_norm.g"].to(dev, torch.float32))
                self.Pn[(2 * l + 1) * d: (2 * l + 2) * d].copy_(base[f"l{l}.mlp_norm.g"].to(dev, torch.float32))
            self.Pn[2 * L * d:].copy_(base["norm.g"].to(dev, torch.float32))
        self.Mo, self.Vo = torch.zeros_like(self.P), torch.zeros_like(self.P)
        self.Mn, self.Vn, self.Gn = torch.zeros_like(self.Pn), torch.zeros_like(self.Pn), torch.zeros_like(self.Pn)
        self.S = torch.empty(self.n_mat, dtype=torch.bfloat16, device=dev)
        self.G = torch.zeros(self.n_mat, dtype=torch.bfloat16, device=dev)
        ops.cast_f32_to_bf16(self.P, self.S)
        self.W = {n: self._view(self.S, n) for n in self.slots}      # bf16 weights the GEMMs read
        self.dW = {n: self._view(self.G, n) for n in self.slots}     # bf16 gradients the GEMMs write
        if D.tie_embeddings:
            self.W["head"], self.dW["head"] = self.W["embed"], self.dW["embed"]
        self.gain = [(self.Pn[2 * l * d: (2 * l + 1) * d], self.Pn[(2 * l + 1) * d: (2 * l + 2) * d]) for l in range(L)]
        self.dgain = [(self.Gn[2 * l * d: (2 * l + 1) * d], self.Gn[(2 * l + 1) * d: (2 * l + 2) * d]) for l in range(L)]
        self.gain_f, self.dgain_f = self.Pn[2 * L * d:], self.Gn[2 * L * d:]
        self.cos, self.sin = rope_tables(D, self.T, dev)
        self.norm_sq = torch.zeros(2, dtype=torch.float32, device=dev)
        self.clip = torch.ones(1, dtype=torch.float32, device=dev)
        self.norm_ws = ops.gradnorm_workspace(self.n_mat, dev)

    def _refresh_shadows(self, cast: bool) -> None:
        if cast:
            ops.cast_f32_to_bf16(self.P, self.S)

    def _export(self, flat: torch.Tensor, norms: torch.Tensor) -> dict:
        """Oracle-named fp32 CPU copies of a flat matrix group + norm group (tests, checkpoints)."""
        D, d, ff, L = self.dims, self.dims.d_model, self.dims.d_ff, self.dims.n_layer
        out = {"embed": self._view(flat, "embed")[: D.vocab].float().cpu()}
        if not D.tie_embeddings:
            out["lm_head.w"] = self._view(flat, "head")[: D.vocab].float().cpu()
        for l in range(L):
            qkv, gu = self._view(flat, f"l{l}.qkv"), self._view(flat, f"l{l}.gu")
            out[f"l{l}.q_proj.w"] = qkv[: self.Nq].float().cpu()
            out[f"l{l}.k_proj.w"] = qkv[self.Nq: self.Nq + self.Nkv].float().cpu()
            out[f"l{l}.v_proj.w"] = qkv[self.Nq + self.Nkv:].float().cpu()
            out[f"l{l}.o_proj.w"] = self._view(flat, f"l{l}.o").float().cpu()
            out[f"l{l}.gate_proj.w"], out[f"l{l}.up_proj.w"] = gu[:ff].float().cpu(), gu[ff:].float().cpu()
            out[f"l{l}.down_proj.w"] = self._view(flat, f"l{l}.down").float().cpu()
            out[f"l{l}.attn_norm.g"] = norms[2 * l * d: (2 * l + 1) * d].float().cpu()
            out[f"l{l}.mlp_norm.g"] = norms[(2 * l + 1) * d: (2 * l + 2) * d].float().cpu()
        out["norm.g"] = norms[2 * L * d:].float().cpu()
        return out

    def params_state(self) -> dict:
        return self._export(self.P, self.Pn)

    def grads_state(self) -> dict:
        return self._export(self.G, self.Gn)

    adapters_state = params_state  # the payload's "trainable tensors" hook

    def grad_views(self, lo: int, hi: int) -> list:
        """Contiguous gradient slices of layers [lo, hi) for the bucketed all-reduce; the top bucket also
        carries an untied lm_head, the bottom one the embedding table and every norm gain."""
        out = [self.G[self.layer_off[lo]: self.layer_off[hi]]]
        if hi == self.dims.n_layer and not self.dims.tie_embeddings:
            out.append(self.dW["head"].view(-1))
        if lo == 0:
            out += [self.dW["embed"].view(-1), self.Gn]
        return out

    def state_parts(self) -> tuple[dict, dict]:
        """(metadata, {name: DEVICE tensor}): 12 bytes per parameter — 96 GB for Llama-3-8B, which is why the payload
        writes the tensors one at a time (train.save_state) instead of building one host dict."""
        meta = {"step": self.step_count, "mode": "full", "n_layer": self.dims.n_layer, "d_model": self.dims.d_model}
        return meta, {"P": self.P, "m": self.Mo, "v": self.Vo, "Pn": self.Pn, "mn": self.Mn, "vn": self.Vn}

    def load_state_dict(self, st: dict) -> None:
        if st.get("mode") != "full" or st["P"].numel() != self.P.numel() or st["Pn"].numel() != self.Pn.numel():
            raise ValueError("checkpoint does not belong to this full-parameter job")
        self.P.copy_(st["P"]); self.Mo.copy_(st["m"]); self.Vo.copy_(st["v"])
        self.Pn.copy_(st["Pn"]); self.Mn.copy_(st["mn"]); self.Vn.copy_(st["vn"])
        self.step_count = int(st["step"])
        self._refresh_shadows(cast=True)

    # ------------------------------------------------------------------ activations
    def _alloc(self) -> None:
        D, dev, M = self.dims, self.dev, self.M
        d, ff, L, Vp = D.d_model, D.d_ff, D.n_layer, D.vocab_padded
        bf, f32 = dict(dtype=torch.bfloat16, device=dev), dict(dtype=torch.float32, device=dev)
        self.x = [torch.empty(M, d, **bf) for _ in range(L + 1)]
        self.x1 = [torch.empty(M, d, **bf) for _ in range(L)]
        self.qkv = [torch.empty(M, self.Nqkv, **bf) for _ in range(L)]
        self.gu = [torch.empty(M, 2 * ff, **bf) for _ in range(L)]
        self.rs1 = [torch.empty(M, **f32) for _ in range(L)]
        self.rs2 = [torch.empty(M, **f32) for _ in range(L)]
        self.rsf, self.rs_tmp = torch.empty(M, **f32), torch.empty(M, **f32)
        self.h, self.act = torch.empty(M, d, **bf), torch.empty(M, ff, **bf)
        self.hf = torch.empty(M, d, **bf)
        self.logits = torch.empty(M, Vp, **bf)
        self.targets = torch.empty(self.B, self.T, dtype=torch.int64, device=dev)
        self.ids = torch.empty(self.B, self.T, dtype=torch.int64, device=dev)
        self.loss = torch.zeros(1, **f32)
        self.xent_ws = ops.softmax_xent_workspace(M, dev)
        self.dxs = [torch.empty(M, d, **bf) for _ in range(3)]
        self.dact, self.dgu = torch.empty(M, ff, **bf), torch.empty(M, 2 * ff, **bf)
        self.dln, self.da = torch.empty(M, d, **bf), torch.empty(M, self.Nq, **bf)
        self.dqkv = torch.empty(M, self.Nqkv, **bf)
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
        return self._attn_out(o)

    def _attn_out(self, o: torch.Tensor) -> torch.Tensor:
        a = o.detach().permute(0, 2, 1, 3)
        if not a.is_contiguous():
            a = a.contiguous()
        return a.view(self.M, self.Nq)

    def _attn_bwd(self, l: int, da: torch.Tensor) -> None:
        """dq | dk | dv land side by side in self.dqkv (the layout of the fused projection's output)."""
        if self.attention == "own":
            self._own_attn_bwd(l, da)
            return
        B, T, D = self.B, self.T, self.dims
        q, k, v, o = self.attn_saved[l]
        go = da.view(B, T, D.n_head, D.d_head).permute(0, 2, 1, 3)
        grads = torch.autograd.grad(o, (q, k, v), go)
        self.attn_saved[l] = None
        c0 = 0
        for g, n, heads in zip(grads, (self.Nq, self.Nkv, self.Nkv), (D.n_head, D.n_kv_head, D.n_kv_head)):
            self.dqkv[:, c0: c0 + n].view(B, T, heads, D.d_head).copy_(g.permute(0, 2, 1, 3))
            c0 += n

    # ------------------------------------------------------------------ the step
    @torch.no_grad()
    def forward_loss(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        D, W = self.dims, self.W
        self.ids.copy_(ids)
        ops.embed_fwd(self.ids.view(-1), W["embed"], None, self.x[0], seq=self.T, vocab=D.vocab)
        for l in range(D.n_layer):
            g_attn, g_mlp = self.gain[l]
            ops.rmsnorm_fwd(self.x[l], g_attn, self.h, self.rs1[l], D.rms_eps)
            ops.gemm(self.h, W[f"l{l}.qkv"], self.qkv[l])
            ops.rope_inplace(self.qkv[l], self.cos, self.sin, seq=self.T, n_heads=D.n_head + D.n_kv_head,
                             d_head=D.d_head)
            a = self._attn_fwd(l)
            ops.gemm(a, W[f"l{l}.o"], self.x1[l], epilogue=EPI_RESIDUAL, aux=self.x[l])
            ops.rmsnorm_fwd(self.x1[l], g_mlp, self.h, self.rs2[l], D.rms_eps)
            ops.gemm(self.h, W[f"l{l}.gu"], self.gu[l])
            ops.swiglu_fwd(self.gu[l], self.act)
            ops.gemm(self.act, W[f"l{l}.down"], self.x[l + 1], epilogue=EPI_RESIDUAL, aux=self.x1[l])
        ops.rmsnorm_fwd(self.x[D.n_layer], self.gain_f, self.hf, self.rsf, D.rms_eps)
        ops.gemm(self.hf, W["head"], self.logits)
        ops.shift_targets(labels, self.targets)
        ops.softmax_xent_fwd_bwd(self.logits, self.targets.view(-1), self.loss, self.xent_ws, vocab=D.vocab)
        return self.loss

    @torch.no_grad()
    def backward(self) -> None:
        D, W, dW = self.dims, self.W, self.dW
        L, mc = D.n_layer, self.bwd_ctas
        if mc:   # SMs are left to an all-reduce on another stream: the next kernel's CTAs must not be parked on them
            ops.set_dependent_launch(False)
        ops.zero(self.Gn)
        dx, dx1, dx0 = self.dxs
        # lm_head: dW = dlogits^T . hf (both MN-major), dhf = dlogits . W_head (B MN-major)
        ops.gemm(self.logits, self.hf, dW["head"], a_mn=True, b_mn=True, max_ctas=mc)
        ops.gemm(self.logits, W["head"], self.dln, b_mn=True, max_ctas=mc)
        ops.rmsnorm_bwd_dgamma(self.dln, self.x[L], self.rsf, self.dgain_f)
        ops.rmsnorm_bwd_dx(self.dln, self.x[L], self.gain_f, self.rsf, dx)
        for l in range(L - 1, -1, -1):
            g_attn, g_mlp = self.gain[l]
            dg_attn, dg_mlp = self.dgain[l]
            # MLP branch (the SwiGLU product and h2 are recomputed, the former inside swiglu_bwd)
            ops.gemm(dx, W[f"l{l}.down"], self.dact, b_mn=True, max_ctas=mc)
            ops.swiglu_bwd(self.dact, self.gu[l], self.dgu, act=self.act)
            ops.gemm(dx, self.act, dW[f"l{l}.down"], a_mn=True, b_mn=True, max_ctas=mc)
            ops.rmsnorm_fwd(self.x1[l], g_mlp, self.h, self.rs_tmp, D.rms_eps)
            ops.gemm(self.dgu, self.h, dW[f"l{l}.gu"], a_mn=True, b_mn=True, max_ctas=mc)
            ops.gemm(self.dgu, W[f"l{l}.gu"], self.dln, b_mn=True, max_ctas=mc)
            ops.rmsnorm_bwd_dgamma(self.dln, self.x1[l], self.rs2[l], dg_mlp)
            ops.rmsnorm_bwd_dx(self.dln, self.x1[l], g_mlp, self.rs2[l], dx1, dres=dx)
            # attention branch
            a = self.att_out[l] if self.attention == "own" else self._attn_out(self.attn_saved[l][3])
            ops.gemm(dx1, a, dW[f"l{l}.o"], a_mn=True, b_mn=True, max_ctas=mc)
            ops.gemm(dx1, W[f"l{l}.o"], self.da, b_mn=True, max_ctas=mc)
            self._attn_bwd(l, self.da)
            ops.rope_inplace(self.dqkv, self.cos, self.sin, seq=self.T, n_heads=D.n_head + D.n_kv_head,
                             d_head=D.d_head, backward=True)
            ops.rmsnorm_fwd(self.x[l], g_attn, self.h, self.rs_tmp, D.rms_eps)
            ops.gemm(self.dqkv, self.h, dW[f"l{l}.qkv"], a_mn=True, b_mn=True, max_ctas=mc)
            ops.gemm(self.dqkv, W[f"l{l}.qkv"], self.dln, b_mn=True, max_ctas=mc)
            ops.rmsnorm_bwd_dgamma(self.dln, self.x[l], self.rs1[l], dg_attn)
            ops.rmsnorm_bwd_dx(self.dln, self.x[l], g_attn, self.rs1[l], dx0, dres=dx1)
            dx, dx0 = dx0, dx
            if l == 0:
                # embedding table: scatter-add of the residual-stream gradient; a tied table already
                # holds the lm_head gradient and accumulates on top of it
                if not D.tie_embeddings:
                    ops.zero(dW["embed"])
                ops.embed_bwd(self.ids.view(-1), dx, dW["embed"], vocab=D.vocab)
            n_pending = len(self.grad_sync.pending) if self.grad_sync is not None else 0
            if self.grad_sync is not None:
                self.grad_sync(self, l)
            if self.overlap_opt:
                self._update_layer(l, self.grad_sync.pending[n_pending:] if self.grad_sync is not None else ())

    def _adam_kw(self, step: int, coef=None) -> dict:
        o = self.opt
        return dict(step=step, lr=o.lr, beta1=o.beta1, beta2=o.beta2, eps=o.eps, weight_decay=o.weight_decay,
                    grad_scale=1.0 / self.world, grad_scale_dev=coef)

    def _update_layer(self, l: int, reductions) -> None:
        """AdamW on layer l's slice of the flat group, on the optimizer stream, one block per SM so that the
        persistent GEMM's CTAs stay resident beside it."""
        lo, hi = self.layer_off[l], self.layer_off[l + 1]
        done = torch.cuda.Event()
        done.record()                                   # every gradient GEMM of the layer has been issued
        with torch.cuda.stream(self.opt_stream):
            self.opt_stream.wait_event(done)
            for work in reductions:                     # and its all-reduce bucket has been summed
                work.wait()
            ops.adamw_step(self.P[lo:hi], self.Mo[lo:hi], self.Vo[lo:hi], self.G[lo:hi], shadow=self.S[lo:hi],
                           max_blocks=ops.NUM_SMS, **self._adam_kw(self.step_count + 1))

    @torch.no_grad()
    def optimizer_step(self) -> None:
        o = self.opt
        self.step_count += 1
        gs = 1.0 / self.world
        coef = None
        if o.max_grad_norm > 0:
            # global norm over both groups: sum of the two squared norms, then one coefficient
            ops.gradnorm_sq(self.G, self.norm_sq[0:1], self.norm_ws)
            ops.gradnorm_sq(self.Gn, self.norm_sq[1:2], self.norm_ws)
            self.norm_sq[0:1].add_(self.norm_sq[1:2])
            ops.clip_coef(self.norm_sq[0:1], self.clip, max_norm=o.max_grad_norm, pre_scale=gs)
            coef = self.clip
        kw = self._adam_kw(self.step_count, coef)
        if self.overlap_opt:
            # the layers were updated during backward; what is left is the embedding table, an untied lm_head
            # and the norm gains
            for lo, hi in ((0, self.layer_off[0]), (self.layer_off[-1], self.n_mat)):
                if hi > lo:
                    ops.adamw_step(self.P[lo:hi], self.Mo[lo:hi], self.Vo[lo:hi], self.G[lo:hi], shadow=self.S[lo:hi], **kw)
            torch.cuda.current_stream().wait_stream(self.opt_stream)
        else:
            ops.adamw_step(self.P, self.Mo, self.Vo, self.G, shadow=self.S, **kw)
        ops.adamw_step(self.Pn, self.Mn, self.Vn, self.Gn, **kw)

    def _optimizer_step_behind_exchange(self) -> bool:
        """Data-parallel step whose buckets are single layers and whose update needs no global norm: instead of waiting
        for the whole exchange and then updating everything, walk the buckets in the order they were sent (top layer
        first), make the stream wait for THAT bucket's all-reduce and update that layer.  What is left of the exchange
        when backward ends (8 x B200, Llama-3-8B: ~20 ms of a 16 GB exchange that backward cannot hide completely)
        then runs under the update of the layers that were reduced long ago.  Returns False when it does not apply."""
        sync = self.grad_sync
        L = self.dims.n_layer
        if (sync is None or self.overlap_opt or self.opt.max_grad_norm > 0 or not hasattr(sync, "buckets")
                or len(sync.buckets) != L or not sync.pending):
            return False
        works = list(sync.pending)
        n_expected = sum(len(self.grad_views(lo, hi)) for lo, hi in sync.buckets)
        if len(works) != n_expected:
            return False
        self.step_count += 1
        kw = self._adam_kw(self.step_count)
        i = 0
        for lo, hi in sync.buckets:                 # issue order = completion order on the communicator's stream
            n = len(self.grad_views(lo, hi))
            for w in works[i: i + n]:
                w.wait()                            # stream-level wait; the host runs ahead
            i += n
            a, b = self.layer_off[lo], self.layer_off[hi]
            ops.adamw_step(self.P[a:b], self.Mo[a:b], self.Vo[a:b], self.G[a:b], shadow=self.S[a:b], **kw)
        sync.pending.clear()
        for a, b in ((0, self.layer_off[0]), (self.layer_off[-1], self.n_mat)):   # embedding table, untied lm_head
            if b > a:
                ops.adamw_step(self.P[a:b], self.Mo[a:b], self.Vo[a:b], self.G[a:b], shadow=self.S[a:b], **kw)
        ops.adamw_step(self.Pn, self.Mn, self.Vn, self.Gn, **kw)
        return True

    def step(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        self.forward_loss(ids, labels)
        self.backward()
        if not self._optimizer_step_behind_exchange():
            if self.grad_sync is not None:
                self.grad_sync(self, -1)
            if self.bwd_ctas:
                ops.set_dependent_launch(True)
            self.optimizer_step()
        elif self.bwd_ctas:
            ops.set_dependent_launch(True)
        return self.loss
