"""ORACLE — test infrastructure, not product code.

CPU restatement (plain PyTorch, fp32, autograd) of the Llama-family LoRA fine-tuning step:
BASELINE.json configs[2..4] (Llama-3.2-1B / Llama-3-8B, LoRA r=16 on q_proj and v_proj).
As for GPT-2 (oracle/gpt2_lora_ref.py) the reference repository contains no compute payload
to restate (reference app/models/examples/mnist.py:41-53 names an external image), so the
step is restated from the published architecture and PINNED against an independent
implementation present in this image: `transformers.models.llama.modeling_llama.
LlamaForCausalLM` (transformers 5.5.0) built from config with this module's weights copied in
— see `hf_crosscheck()` and tests/test_oracle_llama.py.  With respect to the reference
repository itself parity is unpinned (it holds no vectors for this path).

Conventions restated: RMSNorm (x * rsqrt(mean(x^2) + eps) * g), rotary embedding in HF's
rotate_half form with optional "llama3" frequency scaling, grouped-query attention (KV heads
repeated), SwiGLU MLP (down(silu(gate(x)) * up(x))), no biases, nn.Linear weights [out, in],
tied or untied lm_head, HF-shifted cross-entropy.  LoRA (PEFT convention):
y = W x + (alpha/r) B (A x), A [r, in] ~ kaiming_uniform(a=sqrt 5), B [out, r] = 0.

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import torch
import torch.nn.functional as F

from .gpt2_lora_ref import AdamWConfig, causal_lm_loss  # shared pieces of the step


@dataclass(frozen=True)
class LlamaConfig:
    n_layer: int = 16
    d_model: int = 2048
    n_head: int = 32
    n_kv_head: int = 8
    d_ff: int = 8192
    vocab: int = 128256
    ctx: int = 131072
    rms_eps: float = 1e-5
    rope_theta: float = 500000.0
    rope_llama3: bool = True          # Llama-3.1/3.2 frequency scaling
    rope_factor: float = 32.0
    rope_low_freq_factor: float = 1.0
    rope_high_freq_factor: float = 4.0
    rope_original_ctx: int = 8192
    tie_embeddings: bool = True
    d_head_override: int = 0

    @property
    def d_head(self) -> int:
        return self.d_head_override or self.d_model // self.n_head


LLAMA_3_2_1B = LlamaConfig()
LLAMA_3_8B = LlamaConfig(n_layer=32, d_model=4096, n_head=32, n_kv_head=8, d_ff=14336, ctx=8192,
                         rope_llama3=False, tie_embeddings=False)
LLAMA_TINY = LlamaConfig(n_layer=2, d_model=64, n_head=4, n_kv_head=2, d_ff=128, vocab=503, ctx=128,
                         rope_theta=10000.0, rope_llama3=False, tie_embeddings=True)


@dataclass(frozen=True)
class LoRAConfig:
    r: int = 16
    alpha: float = 32.0
    targets: tuple = ("q_proj", "v_proj")

    @property
    def scale(self) -> float:
        return self.alpha / self.r


def init_base_params(cfg: LlamaConfig, seed: int = 0, randomize_all: bool = False) -> dict:
    g = torch.Generator().manual_seed(seed)
    d, ff, dh = cfg.d_model, cfg.d_ff, cfg.d_head

    def nrm(*shape, std=0.02):
        return torch.randn(*shape, generator=g, dtype=torch.float32) * std

    def gain(n):
        return 1.0 + nrm(n, std=0.1) if randomize_all else torch.ones(n)

    p = {"embed": nrm(cfg.vocab, d)}
    for l in range(cfg.n_layer):
        p[f"l{l}.attn_norm.g"] = gain(d)
        p[f"l{l}.q_proj.w"] = nrm(cfg.n_head * dh, d)
        p[f"l{l}.k_proj.w"] = nrm(cfg.n_kv_head * dh, d)
        p[f"l{l}.v_proj.w"] = nrm(cfg.n_kv_head * dh, d)
        p[f"l{l}.o_proj.w"] = nrm(d, cfg.n_head * dh)
        p[f"l{l}.mlp_norm.g"] = gain(d)
        p[f"l{l}.gate_proj.w"] = nrm(ff, d)
        p[f"l{l}.up_proj.w"] = nrm(ff, d)
        p[f"l{l}.down_proj.w"] = nrm(d, ff)
    p["norm.g"] = gain(d)
    if not cfg.tie_embeddings:
        p["lm_head.w"] = nrm(cfg.vocab, d)
    return p


def init_lora_params(cfg: LlamaConfig, lora: LoRAConfig, seed: int = 1, b_init: str = "zero") -> dict:
    g = torch.Generator().manual_seed(seed)
    d, dh = cfg.d_model, cfg.d_head
    outs = {"q_proj": cfg.n_head * dh, "k_proj": cfg.n_kv_head * dh, "v_proj": cfg.n_kv_head * dh, "o_proj": d}
    out = {}
    for l in range(cfg.n_layer):
        for t in lora.targets:
            d_in = cfg.n_head * dh if t == "o_proj" else d
            bound = 1.0 / math.sqrt(d_in)
            out[f"l{l}.{t}.lora_A"] = (torch.rand(lora.r, d_in, generator=g) * 2 - 1) * bound      # [r, in]
            out[f"l{l}.{t}.lora_B"] = (torch.zeros(outs[t], lora.r) if b_init == "zero"
                                       else torch.randn(outs[t], lora.r, generator=g) * 0.02)   # [out, r]
    return out


def synthetic_batch(cfg: LlamaConfig, batch: int, seq: int, seed: int = 1234, rank: int = 0):
    g = torch.Generator().manual_seed(seed + rank)
    ids = torch.randint(0, cfg.vocab, (batch, seq), generator=g, dtype=torch.int64)
    return ids, ids.clone()


# --------------------------------------------------------------------------------------
def rope_inv_freq(cfg: LlamaConfig) -> torch.Tensor:
    """1/theta^(2i/dh), with the Llama-3.1 'llama3' wavelength-dependent rescaling when enabled."""
    dh = cfg.d_head
    inv = 1.0 / (cfg.rope_theta ** (torch.arange(0, dh, 2, dtype=torch.float32) / dh))
    if not cfg.rope_llama3:
        return inv
    low_wl = cfg.rope_original_ctx / cfg.rope_low_freq_factor
    high_wl = cfg.rope_original_ctx / cfg.rope_high_freq_factor
    wl = 2 * math.pi / inv
    scaled = torch.where(wl > low_wl, inv / cfg.rope_factor, inv)
    smooth = (cfg.rope_original_ctx / wl - cfg.rope_low_freq_factor) / (cfg.rope_high_freq_factor - cfg.rope_low_freq_factor)
    mid = (1 - smooth) * inv / cfg.rope_factor + smooth * inv
    is_mid = (wl <= low_wl) & (wl >= high_wl)
    return torch.where(is_mid, mid, scaled)


def rope_tables(cfg: LlamaConfig, seq: int):
    """cos, sin [seq, dh/2] (position t, frequency i)."""
    ang = torch.arange(seq, dtype=torch.float32)[:, None] * rope_inv_freq(cfg)[None, :]
    return ang.cos(), ang.sin()


def apply_rope(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor) -> torch.Tensor:
    """x [B, H, T, dh]; HF rotate_half convention: pairs are (i, i + dh/2)."""
    half = x.shape[-1] // 2
    x1, x2 = x[..., :half], x[..., half:]
    c, s = cos[None, None], sin[None, None]
    return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1)


def rmsnorm(x: torch.Tensor, g: torch.Tensor, eps: float) -> torch.Tensor:
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * g


def _linear(x, p, lp, name, lora: LoRAConfig | None):
    y = x @ p[name + ".w"].t()
    if lora is not None and lp is not None and (name + ".lora_A") in lp:
        y = y + lora.scale * ((x @ lp[name + ".lora_A"].t()) @ lp[name + ".lora_B"].t())
    return y


def forward(p: dict, lp: dict | None, ids: torch.Tensor, cfg: LlamaConfig, lora: LoRAConfig | None) -> torch.Tensor:
    B, T = ids.shape
    H, KV, dh = cfg.n_head, cfg.n_kv_head, cfg.d_head
    cos, sin = rope_tables(cfg, T)
    x = F.embedding(ids, p["embed"])
    mask = torch.ones(T, T, dtype=torch.bool).tril()
    for l in range(cfg.n_layer):
        h = rmsnorm(x, p[f"l{l}.attn_norm.g"], cfg.rms_eps)
        q = _linear(h, p, lp, f"l{l}.q_proj", lora).view(B, T, H, dh).transpose(1, 2)
        k = _linear(h, p, lp, f"l{l}.k_proj", lora).view(B, T, KV, dh).transpose(1, 2)
        v = _linear(h, p, lp, f"l{l}.v_proj", lora).view(B, T, KV, dh).transpose(1, 2)
        q, k = apply_rope(q, cos, sin), apply_rope(k, cos, sin)
        k = k.repeat_interleave(H // KV, dim=1)
        v = v.repeat_interleave(H // KV, dim=1)
        att = (q @ k.transpose(-1, -2)) / math.sqrt(dh)
        att = torch.softmax(att.masked_fill(~mask, float("-inf")), dim=-1)
        a = (att @ v).transpose(1, 2).reshape(B, T, H * dh)
        x = x + _linear(a, p, lp, f"l{l}.o_proj", lora)
        h = rmsnorm(x, p[f"l{l}.mlp_norm.g"], cfg.rms_eps)
        act = F.silu(_linear(h, p, lp, f"l{l}.gate_proj", lora)) * _linear(h, p, lp, f"l{l}.up_proj", lora)
        x = x + _linear(act, p, lp, f"l{l}.down_proj", lora)
    x = rmsnorm(x, p["norm.g"], cfg.rms_eps)
    head = p["embed"] if cfg.tie_embeddings else p["lm_head.w"]
    return x @ head.t()


class OracleTrainer:
    """fwd + shifted CE + autograd bwd + torch.optim.AdamW(foreach=False, fused=False) on the adapters."""

    def __init__(self, cfg: LlamaConfig, lora: LoRAConfig, opt: AdamWConfig, base: dict, adapters: dict):
        self.cfg, self.lora, self.opt_cfg = cfg, lora, opt
        self.base = {k: v.clone().float() for k, v in base.items()}
        self.trainable = {k: v.clone().float().requires_grad_(True) for k, v in adapters.items()}
        self.optim = torch.optim.AdamW(list(self.trainable.values()), lr=opt.lr, betas=(opt.beta1, opt.beta2),
                                       eps=opt.eps, weight_decay=opt.weight_decay, foreach=False, fused=False)
        self.last_grads: dict = {}

    def step(self, ids, labels) -> float:
        self.optim.zero_grad(set_to_none=True)
        loss = causal_lm_loss(forward(self.base, self.trainable, ids, self.cfg, self.lora), labels)
        loss.backward()
        if self.opt_cfg.max_grad_norm > 0:
            torch.nn.utils.clip_grad_norm_(list(self.trainable.values()), self.opt_cfg.max_grad_norm)
        self.last_grads = {k: t.grad.detach().clone() for k, t in self.trainable.items()}
        self.optim.step()
        return float(loss.detach())


class FullOracleTrainer:
    """Full-parameter fine-tune (BASELINE.json configs[4]): every tensor of `base` is trainable, no
    adapters; fwd + shifted CE + autograd bwd + torch.optim.AdamW(foreach=False, fused=False) with one
    uniform weight decay over all tensors."""

    def __init__(self, cfg: LlamaConfig, opt: AdamWConfig, base: dict):
        self.cfg, self.opt_cfg = cfg, opt
        self.trainable = {k: v.clone().float().requires_grad_(True) for k, v in base.items()}
        self.optim = torch.optim.AdamW(list(self.trainable.values()), lr=opt.lr, betas=(opt.beta1, opt.beta2),
                                       eps=opt.eps, weight_decay=opt.weight_decay, foreach=False, fused=False)
        self.last_grads: dict = {}

    def grads(self, ids, labels):
        for t in self.trainable.values():
            t.grad = None
        loss = causal_lm_loss(forward(self.trainable, None, ids, self.cfg, None), labels)
        loss.backward()
        return float(loss.detach()), {k: t.grad.detach().clone() for k, t in self.trainable.items()}

    def step(self, ids, labels, world_grads: list | None = None) -> float:
        """One optimizer step.  `world_grads` = [(ids, labels) per rank] emulates data parallelism: the
        gradient is the mean of the per-rank gradients and the returned loss the mean of the per-rank
        losses (same convention as gpt2_lora_ref.OracleTrainer.step)."""
        batches = world_grads if world_grads is not None else [(ids, labels)]
        losses, per_rank = [], []
        for b_ids, b_labels in batches:
            loss, g = self.grads(b_ids, b_labels)
            losses.append(loss); per_rank.append(g)
        for k, t in self.trainable.items():
            t.grad = sum(g[k] for g in per_rank) / len(per_rank)
        if self.opt_cfg.max_grad_norm > 0:
            torch.nn.utils.clip_grad_norm_(list(self.trainable.values()), self.opt_cfg.max_grad_norm)
        self.last_grads = {k: t.grad.detach().clone() for k, t in self.trainable.items()}
        self.optim.step()
        return sum(losses) / len(losses)


def hf_model(cfg: LlamaConfig, p: dict):
    """transformers' LlamaForCausalLM built from config holding this module's parameters `p`."""
    from transformers import LlamaConfig as HFConfig, LlamaForCausalLM

    rope = {"rope_type": "default", "rope_theta": cfg.rope_theta}
    if cfg.rope_llama3:
        rope = {"rope_type": "llama3", "rope_theta": cfg.rope_theta, "factor": cfg.rope_factor,
                "low_freq_factor": cfg.rope_low_freq_factor, "high_freq_factor": cfg.rope_high_freq_factor,
                "original_max_position_embeddings": cfg.rope_original_ctx}
    hf_cfg = HFConfig(vocab_size=cfg.vocab, hidden_size=cfg.d_model, intermediate_size=cfg.d_ff,
                      num_hidden_layers=cfg.n_layer, num_attention_heads=cfg.n_head,
                      num_key_value_heads=cfg.n_kv_head, head_dim=cfg.d_head, max_position_embeddings=cfg.ctx,
                      rms_norm_eps=cfg.rms_eps, rope_parameters=rope, tie_word_embeddings=cfg.tie_embeddings,
                      attention_bias=False, mlp_bias=False, attention_dropout=0.0, attn_implementation="eager")
    hf = LlamaForCausalLM(hf_cfg).float().eval()
    sd = hf.state_dict()
    with torch.no_grad():
        sd["model.embed_tokens.weight"].copy_(p["embed"])
        for l in range(cfg.n_layer):
            pre = f"model.layers.{l}."
            sd[pre + "input_layernorm.weight"].copy_(p[f"l{l}.attn_norm.g"])
            sd[pre + "post_attention_layernorm.weight"].copy_(p[f"l{l}.mlp_norm.g"])
            for n in ("q_proj", "k_proj", "v_proj", "o_proj"):
                sd[pre + f"self_attn.{n}.weight"].copy_(p[f"l{l}.{n}.w"])
            for n in ("gate_proj", "up_proj", "down_proj"):
                sd[pre + f"mlp.{n}.weight"].copy_(p[f"l{l}.{n}.w"])
        sd["model.norm.weight"].copy_(p["norm.g"])
        if not cfg.tie_embeddings:
            sd["lm_head.weight"].copy_(p["lm_head.w"])
    if cfg.tie_embeddings:
        hf.tie_weights()
    return hf


def hf_crosscheck(cfg: LlamaConfig = LLAMA_TINY, batch: int = 2, seq: int = 24, seed: int = 0) -> dict:
    """transformers' LlamaForCausalLM with this module's weights: logits / loss / grads."""
    p = init_base_params(cfg, seed, randomize_all=True)
    hf = hf_model(cfg, p)
    ids, labels = synthetic_batch(cfg, batch, seq)
    mine = {k: v.clone().requires_grad_(True) for k, v in p.items()}
    logits = forward(mine, None, ids, cfg, None)
    loss = causal_lm_loss(logits, labels)
    loss.backward()
    out = hf(input_ids=ids, labels=labels)
    out.loss.backward()
    return {
        "logits_max_abs_diff": float((logits.detach() - out.logits.detach()).abs().max()),
        "loss_mine": float(loss.detach()), "loss_hf": float(out.loss.detach()),
        "q0_grad_max_abs_diff": float((mine["l0.q_proj.w"].grad - hf.model.layers[0].self_attn.q_proj.weight.grad).abs().max()),
        "down0_grad_max_abs_diff": float((mine["l0.down_proj.w"].grad - hf.model.layers[0].mlp.down_proj.weight.grad).abs().max()),
    }


if __name__ == "__main__":  # pragma: no cover
    print(hf_crosscheck())
    print(hf_crosscheck(LlamaConfig(n_layer=1, d_model=128, n_head=4, n_kv_head=2, d_ff=256, vocab=1000, ctx=16384,
                                    rope_llama3=True, rope_original_ctx=64), batch=1, seq=96))
