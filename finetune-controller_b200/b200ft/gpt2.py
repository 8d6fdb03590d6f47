"""GPT-2 LoRA fine-tuning step on one B200, written against the C ABI (b200ft.ops).

This is the product path of SURVEY.md §8(a2): every GEMM is the tcgen05 kernel (the LoRA
branch rides in as a second K segment of the frozen projection, forward and backward), the
loss is the fused softmax-cross-entropy, the update is the fused AdamW.  Activations are bf16
row-major [tokens, features]; statistics, accumulators and optimizer state are fp32.

Frozen weights are kept in HBM in the two layouts the K-major GEMM wants: W^T [out, in] for
forward (y = x W) and W [in, out] (the HF Conv1D layout) for backward (dx = dy W^T).

Attention: `attention="own"` runs this library's tcgen05 flash-attention kernels
(attention.py), leaving no library kernel in the step — what "auto", the default, selects for 64- and 128-wide
heads; `attention="sdpa"` delegates it to torch's scaled_dot_product_attention -> cuDNN, forward and backward
(library; kept for A/B measurements and for other head sizes).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn.functional as F

from . import ops
from .attention import OwnAttentionMixin, resolve_attention
from ._lib import EPI_DGELU, EPI_GELU, EPI_NONE, EPI_RESIDUAL

R_PAD = 16  # adapter rank padded to one UMMA K step (16 bf16)


@dataclass(frozen=True)
class GPT2Dims:
    n_layer: int = 12
    d_model: int = 768
    n_head: int = 12
    d_ff: int = 3072
    vocab: int = 50257
    ctx: int = 1024
    ln_eps: float = 1e-5

    @property
    def vocab_padded(self) -> int:
        return (self.vocab + 127) // 128 * 128


@dataclass(frozen=True)
class LoRASpec:
    r: int = 8
    alpha: float = 16.0

    @property
    def scale(self) -> float:
        return self.alpha / self.r


@dataclass(frozen=True)
class OptSpec:
    lr: float = 1e-4
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 0.01
    max_grad_norm: float = 0.0


class StepHostMixin:
    """What every step engine offers to the payload besides step(): the host-buffer call and the
    resume state.  Engines provide B, T, dev, loss, step(), P / Mo / Vo, step_count, lora, dims,
    _refresh_shadows()."""

    def step_host(self, ids_host: torch.Tensor, labels_host: torch.Tensor) -> float:
        """The call a user of the package makes: HOST int64 [batch, seq] tensors in (pinned
        memory makes the copies asynchronous), python float loss out.  Copies the inputs to
        the device, runs one full step, reads the loss back."""
        if not hasattr(self, "_ids_dev"):
            self._ids_dev = torch.empty(self.B, self.T, dtype=torch.int64, device=self.dev)
            self._labels_dev = torch.empty(self.B, self.T, dtype=torch.int64, device=self.dev)
            self._loss_host = torch.empty(1, dtype=torch.float32).pin_memory()
        self._ids_dev.copy_(ids_host, non_blocking=True)
        self._labels_dev.copy_(labels_host, non_blocking=True)
        self.step(self._ids_dev, self._labels_dev)
        self._loss_host.copy_(self.loss, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        return float(self._loss_host[0])

    def state_parts(self) -> tuple[dict, dict]:
        """(small metadata dict, {name: DEVICE tensor}) — everything needed to continue training.  The payload writes
        the tensors one at a time, so the host never holds more than one of them (train.save_state)."""
        meta = {"step": self.step_count, "r": self.lora.r, "alpha": self.lora.alpha, "n_layer": self.dims.n_layer,
                "d_model": self.dims.d_model}
        return meta, {"P": self.P, "m": self.Mo, "v": self.Vo}

    def state_dict(self) -> dict:
        """state_parts() as one dict of CPU tensors (tests, small models)."""
        meta, tensors = self.state_parts()
        return {**meta, **{k: t.detach().cpu() for k, t in tensors.items()}}

    def load_state_dict(self, st: dict) -> None:
        want = (self.lora.r, self.dims.n_layer, self.dims.d_model)
        got = (st["r"], st["n_layer"], st["d_model"])
        if want != got or st["P"].numel() != self.P.numel():
            raise ValueError(f"checkpoint is for (r, layers, d_model) = {got}, this job is {want}")
        self.P.copy_(st["P"]); self.Mo.copy_(st["m"]); self.Vo.copy_(st["v"])
        self.step_count = int(st["step"])
        self._refresh_shadows(cast=True)


class GPT2LoRAStep(StepHostMixin, OwnAttentionMixin):
    """Holds device state for one rank and runs fwd + loss + bwd + AdamW per `step()`."""

    def __init__(self, dims: GPT2Dims, lora: LoRASpec, opt: OptSpec, base: dict, adapters: dict, *,
                 batch: int, seq: int, device="cuda:0", world_size: int = 1, grad_sync=None, attention: str = "auto"):
        if not torch.cuda.is_available():
            raise RuntimeError("GPT2LoRAStep needs a CUDA device: there is no CPU path")
        self.attention = resolve_attention(attention, dims.d_model // dims.n_head)
        assert lora.r <= R_PAD and seq <= dims.ctx
        assert dims.d_model % 8 == 0 and dims.d_ff % 8 == 0
        self.dims, self.lora, self.opt = dims, lora, opt
        self.B, self.T, self.M = batch, seq, batch * seq
        self.dev = torch.device(device)
        self.world = world_size
        self.grad_sync = grad_sync  # callable(engine, layer_lo, layer_hi) or None
        self.step_count = 0
        self._upload_base(base)
        self._init_adapters(adapters)
        self._alloc_activations()

    # ------------------------------------------------------------------ parameters
    def _bf16(self, t: torch.Tensor) -> torch.Tensor:
        src = t.to(self.dev, torch.float32).contiguous()
        out = torch.empty(src.shape, dtype=torch.bfloat16, device=self.dev)
        ops.cast_f32_to_bf16(src, out)
        return out

    def _f32(self, t: torch.Tensor) -> torch.Tensor:
        return t.to(self.dev, torch.float32).contiguous()

    def _upload_base(self, p: dict) -> None:
        D, dev = self.dims, self.dev
        d, Vp = D.d_model, D.vocab_padded
        wte = torch.zeros(Vp, d, dtype=torch.bfloat16, device=dev)
        wte[: D.vocab] = self._bf16(p["wte"])
        self.wte = wte                                              # [Vp, d]  lm_head fwd: B operand
        self.wteT = torch.empty(d, Vp, dtype=torch.bfloat16, device=dev)
        ops.transpose(wte, self.wteT)                               # [d, Vp]  lm_head bwd: B operand
        self.wpe = self._bf16(p["wpe"])
        self.layers = []
        for l in range(D.n_layer):
            L = {}
            for name in ("c_attn", "attn_proj", "c_fc", "mlp_proj"):
                w = self._bf16(p[f"h{l}.{name}.w"])                 # HF Conv1D layout [in, out]
                wt = torch.empty(w.shape[1], w.shape[0], dtype=torch.bfloat16, device=dev)
                ops.transpose(w, wt)                                # [out, in]
                L[name + ".w"], L[name + ".wT"] = w, wt
                L[name + ".b"] = self._f32(p[f"h{l}.{name}.b"])
            for ln in ("ln1", "ln2"):
                L[ln + ".g"], L[ln + ".b"] = self._f32(p[f"h{l}.{ln}.g"]), self._f32(p[f"h{l}.{ln}.b"])
            self.layers.append(L)
        self.lnf_g, self.lnf_b = self._f32(p["lnf.g"]), self._f32(p["lnf.b"])

    def _init_adapters(self, ad: dict) -> None:
        """Flat fp32 master [At_all (L*16, d) | B_all (L*16, 3d)] with zero padding rows.
        At = lora_A^T (so the token contraction lora_grad() emits lands in the master's layout)."""
        D, dev = self.dims, self.dev
        L, d, r = D.n_layer, D.d_model, self.lora.r
        nA, nB = L * R_PAD * d, L * R_PAD * 3 * d
        self.P = torch.zeros(nA + nB, dtype=torch.float32, device=dev)
        self.At_all = self.P[:nA].view(L * R_PAD, d)
        self.B_all = self.P[nA:].view(L * R_PAD, 3 * d)
        for l in range(L):
            self.At_all[l * R_PAD: l * R_PAD + r] = self._f32(ad[f"h{l}.c_attn.lora_A"]).t()
            self.B_all[l * R_PAD: l * R_PAD + r] = self._f32(ad[f"h{l}.c_attn.lora_B"])
        self.Mo = torch.zeros_like(self.P)
        self.Vo = torch.zeros_like(self.P)
        self.G = torch.zeros_like(self.P)
        self.gAt_all = self.G[:nA].view(L * R_PAD, d)
        self.gB_all = self.G[nA:].view(L * R_PAD, 3 * d)
        self.S = torch.empty(nA + nB, dtype=torch.bfloat16, device=dev)   # bf16 shadow of P
        self.At_sh = self.S[:nA].view(L * R_PAD, d)
        self.B_sh = self.S[nA:].view(L * R_PAD, 3 * d)
        self.A_sh = torch.empty(d, L * R_PAD, dtype=torch.bfloat16, device=dev)       # transposes of
        self.Bt_sh = torch.empty(3 * d, L * R_PAD, dtype=torch.bfloat16, device=dev)  # the shadows
        self.norm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.clip = torch.ones(1, dtype=torch.float32, device=dev)
        self.norm_ws = ops.gradnorm_workspace(self.P.numel(), dev)
        self._refresh_shadows(cast=True)

    def _refresh_shadows(self, cast: bool) -> None:
        if cast:
            ops.cast_f32_to_bf16(self.P, self.S)
        ops.transpose(self.At_sh, self.A_sh)
        ops.transpose(self.B_sh, self.Bt_sh)

    def adapters_state(self) -> dict:
        """Adapters in the oracle's naming / layout (lora_A [d, r], lora_B [r, 3d]), fp32, on CPU."""
        r = self.lora.r
        out = {}
        for l in range(self.dims.n_layer):
            out[f"h{l}.c_attn.lora_A"] = self.At_all[l * R_PAD: l * R_PAD + r].t().contiguous().cpu()
            out[f"h{l}.c_attn.lora_B"] = self.B_all[l * R_PAD: l * R_PAD + r].contiguous().cpu()
        return out

    def grads_state(self) -> dict:
        r = self.lora.r
        out = {}
        for l in range(self.dims.n_layer):
            out[f"h{l}.c_attn.lora_A"] = self.gAt_all[l * R_PAD: l * R_PAD + r].t().contiguous().cpu()
            out[f"h{l}.c_attn.lora_B"] = self.gB_all[l * R_PAD: l * R_PAD + r].contiguous().cpu()
        return out

    def grad_views(self, lo: int, hi: int) -> list:
        """Contiguous gradient slices of layers [lo, hi) (what one all-reduce bucket covers).  The flat buffer is laid
        out [A^T of all layers | B of all layers], so a layer range is two slices; the whole model is the one flat buffer
        (1.2 MB for GPT-2 124M: one latency-bound all-reduce per step, which is what the payload and bench use)."""
        if lo == 0 and hi == self.dims.n_layer:
            return [self.G]
        return [self.gAt_all[lo * R_PAD: hi * R_PAD], self.gB_all[lo * R_PAD: hi * R_PAD]]

    # ------------------------------------------------------------------ activations
    def _alloc_activations(self) -> None:
        D, dev, M = self.dims, self.dev, self.M
        d, ff, L, Vp = D.d_model, D.d_ff, D.n_layer, D.vocab_padded
        bf = dict(dtype=torch.bfloat16, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        self.x = [torch.empty(M, d, **bf) for _ in range(L + 1)]     # residual stream at layer inputs
        self.x1 = [torch.empty(M, d, **bf) for _ in range(L)]        # after the attention branch
        self.h1 = [torch.empty(M, d, **bf) for _ in range(L)]        # LN1 output (needed for dA)
        self.U = [torch.empty(M, R_PAD, **bf) for _ in range(L)]     # s * h1 * A
        self.qkv = [torch.empty(M, 3 * d, **bf) for _ in range(L)]
        self.hpre = [torch.empty(M, ff, **bf) for _ in range(L)]     # c_fc output before GELU
        self.st1 = [(torch.empty(M, **f32), torch.empty(M, **f32)) for _ in range(L)]
        self.st2 = [(torch.empty(M, **f32), torch.empty(M, **f32)) for _ in range(L)]
        self.stf = (torch.empty(M, **f32), torch.empty(M, **f32))
        self.h2 = torch.empty(M, d, **bf)
        self.u = torch.empty(M, ff, **bf)
        self.hf = torch.empty(M, d, **bf)
        self.logits = torch.empty(M, Vp, **bf)
        self.targets = torch.empty(self.B, self.T, dtype=torch.int64, device=dev)
        self.loss = torch.zeros(1, **f32)
        self.xent_ws = ops.softmax_xent_workspace(M, dev)
        self.dxs = [torch.empty(M, d, **bf) for _ in range(3)]
        self.dh = torch.empty(M, ff, **bf)
        self.dln = torch.empty(M, d, **bf)
        self.da = torch.empty(M, d, **bf)
        self.dU = torch.empty(M, R_PAD, **bf)
        self.attn_saved = [None] * L
        if self.attention == "own":
            self._own_attn_setup(D.n_head, D.n_head, d // D.n_head, L)

    # ------------------------------------------------------------------ attention: torch SDPA (library) or own kernels
    def _attn_fwd(self, l: int) -> torch.Tensor:
        if self.attention == "own":
            return self._own_attn_fwd(l)
        B, T, H = self.B, self.T, self.dims.n_head
        dh = self.dims.d_model // H
        v5 = self.qkv[l].view(B, T, 3, H, dh)
        q, k, v = (v5[:, :, i].permute(0, 2, 1, 3).detach().requires_grad_(True) for i in range(3))
        with torch.enable_grad():
            o = F.scaled_dot_product_attention(q, k, v, is_causal=True)   # [B,H,T,dh]
        self.attn_saved[l] = (q, k, v, o)
        a = o.detach().permute(0, 2, 1, 3)
        if not a.is_contiguous():
            a = a.contiguous()
        return a.view(self.M, self.dims.d_model)

    def _attn_bwd(self, l: int, da: torch.Tensor):
        """Returns (dq, dk, dv), each a [tokens, d_model] matrix.  The library hands them back as
        three separate dense buffers; they are consumed as separate K segments downstream, so
        nothing is copied into an interleaved [tokens, 3*d_model] layout.  (The own kernels write one
        [tokens, 3*d_model] buffer; its column blocks are returned.)"""
        if self.attention == "own":
            return self._own_attn_bwd(l, da)
        B, T, H = self.B, self.T, self.dims.n_head
        dh = self.dims.d_model // H
        q, k, v, o = self.attn_saved[l]
        go = da.view(B, T, H, dh).permute(0, 2, 1, 3)
        grads = torch.autograd.grad(o, (q, k, v), go)
        self.attn_saved[l] = None
        out = []
        for g in grads:
            g = g.permute(0, 2, 1, 3)              # [B, T, H, dh]
            if not g.is_contiguous():
                g = g.contiguous()
            out.append(g.view(self.M, self.dims.d_model))
        return out

    # ------------------------------------------------------------------ the step
    @torch.no_grad()
    def forward_loss(self, ids: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        D, M = self.dims, self.M
        d, s = D.d_model, self.lora.scale
        ops.embed_fwd(ids.view(-1), self.wte, self.wpe, self.x[0], seq=self.T, vocab=D.vocab)
        for l, W in enumerate(self.layers):
            rows = slice(l * R_PAD, (l + 1) * R_PAD)
            ops.layernorm_fwd(self.x[l], W["ln1.g"], W["ln1.b"], self.h1[l], *self.st1[l], D.ln_eps)
            ops.gemm(self.h1[l], self.At_sh[rows], self.U[l], alpha=s)
            ops.gemm(self.h1[l], W["c_attn.wT"], self.qkv[l], a2=self.U[l], b2=self.Bt_sh[:, rows],
                     bias=W["c_attn.b"])
            a = self._attn_fwd(l)
            ops.gemm(a, W["attn_proj.wT"], self.x1[l], bias=W["attn_proj.b"], epilogue=EPI_RESIDUAL,
                     aux=self.x[l])
            ops.layernorm_fwd(self.x1[l], W["ln2.g"], W["ln2.b"], self.h2, *self.st2[l], D.ln_eps)
            ops.gemm(self.h2, W["c_fc.wT"], self.hpre[l], bias=W["c_fc.b"], epilogue=EPI_GELU, out2=self.u)
            ops.gemm(self.u, W["mlp_proj.wT"], self.x[l + 1], bias=W["mlp_proj.b"], epilogue=EPI_RESIDUAL,
                     aux=self.x1[l])
        ops.layernorm_fwd(self.x[D.n_layer], self.lnf_g, self.lnf_b, self.hf, *self.stf, D.ln_eps)
        ops.gemm(self.hf, self.wte, self.logits)
        ops.shift_targets(labels, self.targets)
        ops.softmax_xent_fwd_bwd(self.logits, self.targets.view(-1), self.loss, self.xent_ws, vocab=D.vocab)
        return self.loss

    @torch.no_grad()
    def backward(self) -> None:
        """Consumes d(loss)/d(logits) left in self.logits by forward_loss(); fills self.G."""
        D = self.dims
        s = self.lora.scale
        ops.zero(self.G)
        dx, dx1, dx0 = self.dxs
        ops.gemm(self.logits, self.wteT, self.dln)                      # d hf = dlogits . wte
        ops.layernorm_bwd_dx(self.dln, self.x[D.n_layer], self.lnf_g, *self.stf, dx)
        for l in range(D.n_layer - 1, -1, -1):
            W = self.layers[l]
            rows = slice(l * R_PAD, (l + 1) * R_PAD)
            # MLP branch
            ops.gemm(dx, W["mlp_proj.w"], self.dh, epilogue=EPI_DGELU, aux=self.hpre[l])
            ops.gemm(self.dh, W["c_fc.w"], self.dln)
            ops.layernorm_bwd_dx(self.dln, self.x1[l], W["ln2.g"], *self.st2[l], dx1, dres=dx)
            # attention branch
            ops.gemm(dx1, W["attn_proj.w"], self.da)
            dq, dk, dv = self._attn_bwd(l, self.da)
            # adapter gradients (dB = U^T dqkv, dU = s dqkv B^T, dA^T = dU^T h1) and d ln1
            d = D.d_model
            gB, Bsh, Wqkv = self.gB_all[rows], self.B_sh[rows], W["c_attn.w"]
            if self.attention == "own":
                # the own kernels leave dq | dk | dv side by side in one [tokens, 3d] buffer: one contraction each
                dqkv = self.dqkv
                ops.lora_grad(self.U[l], dqkv, gB, r=R_PAD, cols=3 * d)
                ops.gemm(dqkv, Bsh, self.dU, alpha=s)
                ops.lora_grad(self.dU, self.h1[l], self.gAt_all[rows], r=R_PAD, cols=d)
                ops.gemm(dqkv, Wqkv, self.dln, a2=self.dU, b2=self.A_sh[:, rows])
            else:
                # the library hands back three separate buffers: each third is its own K segment / column block
                for i, g in enumerate((dq, dk, dv)):
                    ops.lora_grad(self.U[l], g, gB[:, i * d:(i + 1) * d], r=R_PAD, cols=d)
                ops.gemm(dq, Bsh[:, :d], self.dU, a2=dk, b2=Bsh[:, d:2 * d], more=[(dv, Bsh[:, 2 * d:])], alpha=s)
                ops.lora_grad(self.dU, self.h1[l], self.gAt_all[rows], r=R_PAD, cols=d)
                ops.gemm(dq, Wqkv[:, :d], self.dln, a2=dk, b2=Wqkv[:, d:2 * d],
                         more=[(dv, Wqkv[:, 2 * d:]), (self.dU, self.A_sh[:, rows])])
            ops.layernorm_bwd_dx(self.dln, self.x[l], W["ln1.g"], *self.st1[l], dx0, dres=dx1)
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
        """ids, labels: int64 [batch, seq] on the device.  Returns the (device) loss scalar."""
        self.forward_loss(ids, labels)
        self.backward()
        if self.grad_sync is not None:
            self.grad_sync(self, -1)  # flush: wait for outstanding reductions
        self.optimizer_step()
        return self.loss
