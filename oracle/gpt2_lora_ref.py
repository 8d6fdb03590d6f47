"""ORACLE — test infrastructure, not product code.

CPU restatement (plain PyTorch, fp32, autograd) of the fine-tuning step that the
controller's PyTorchJob launches.  The reference repository ships no compute payload
(its only job definition, reference `app/models/examples/mnist.py:41-53`, names an
external image whose source is absent), so `BASELINE.json.north_star` makes the
GPT-2-class LoRA step the reference workload and this file is its restatement
("reference torch CPU path", BASELINE.json configs[0]).

PARITY PINNING: the reference holds no golden vectors for this path (reference
`tests/test_security.py:1-36` and `tests/test_introspection.py:1-48` cover auth only).
The restatement is therefore pinned against an independent implementation that IS
available in this image: `transformers.models.gpt2.modeling_gpt2.GPT2LMHeadModel`
(transformers 5.5.0) with weights copied across — see `hf_crosscheck()` and
`tests/test_oracle.py`.  LoRA follows the published PEFT convention (PEFT itself is not
installed): y = base(x) + (alpha/r) * (x @ A) @ B, A ~ kaiming_uniform(a=sqrt(5)), B = 0.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / --impl reference
legs may import this module.  The product path (`finetune-controller_b200/`) never does.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------------------
# configuration
# --------------------------------------------------------------------------------------
@dataclass(frozen=True)
class GPT2Config:
    n_layer: int = 12
    d_model: int = 768
    n_head: int = 12
    d_ff: int = 3072
    vocab: int = 50257
    ctx: int = 1024
    ln_eps: float = 1e-5

    @property
    def d_head(self) -> int:
        return self.d_model // self.n_head


GPT2_124M = GPT2Config()
GPT2_TINY = GPT2Config(n_layer=2, d_model=64, n_head=2, d_ff=256, vocab=503, ctx=64)


@dataclass(frozen=True)
class LoRAConfig:
    r: int = 8
    alpha: float = 16.0
    # GPT-2: PEFT's default target for gpt2 is the fused qkv projection `c_attn`
    targets: tuple = ("c_attn",)

    @property
    def scale(self) -> float:
        return self.alpha / self.r


@dataclass(frozen=True)
class AdamWConfig:
    lr: float = 1e-4
    beta1: float = 0.9
    beta2: float = 0.999
    eps: float = 1e-8
    weight_decay: float = 0.01
    max_grad_norm: float = 0.0  # 0 = no clipping


# --------------------------------------------------------------------------------------
# deterministic parameter construction (shared with the GPU path through the *values*,
# never through code: the GPU host code receives these tensors from the test harness)
# --------------------------------------------------------------------------------------
def init_base_params(cfg: GPT2Config, seed: int = 0, randomize_all: bool = False) -> dict:
    """GPT-2 parameter set.  Weight layout follows HF GPT-2: Conv1D weights are [in, out].

    randomize_all=False: HF-style init (normal(0, 0.02) weights, zero biases, LN gamma=1
    beta=0).  randomize_all=True additionally draws biases / LN params at random so that a
    parity test cannot pass with a kernel that drops them.
    """
    g = torch.Generator().manual_seed(seed)
    d, ff, L = cfg.d_model, cfg.d_ff, cfg.n_layer

    def nrm(*shape, std=0.02):
        return torch.randn(*shape, generator=g, dtype=torch.float32) * std

    def bias(n):
        return nrm(n) if randomize_all else torch.zeros(n)

    def gamma(n):
        return 1.0 + nrm(n, std=0.1) if randomize_all else torch.ones(n)

    p = {"wte": nrm(cfg.vocab, d), "wpe": nrm(cfg.ctx, d)}
    for l in range(L):
        p[f"h{l}.ln1.g"], p[f"h{l}.ln1.b"] = gamma(d), bias(d)
        p[f"h{l}.c_attn.w"], p[f"h{l}.c_attn.b"] = nrm(d, 3 * d), bias(3 * d)
        p[f"h{l}.attn_proj.w"], p[f"h{l}.attn_proj.b"] = nrm(d, d), bias(d)
        p[f"h{l}.ln2.g"], p[f"h{l}.ln2.b"] = gamma(d), bias(d)
        p[f"h{l}.c_fc.w"], p[f"h{l}.c_fc.b"] = nrm(d, ff), bias(ff)
        p[f"h{l}.mlp_proj.w"], p[f"h{l}.mlp_proj.b"] = nrm(ff, d), bias(d)
    p["lnf.g"], p["lnf.b"] = gamma(d), bias(d)
    return p


def init_lora_params(cfg: GPT2Config, lora: LoRAConfig, seed: int = 1, b_init: str = "zero") -> dict:
    """A: [d_in, r] kaiming_uniform(a=sqrt(5)) over fan_in=d_in (PEFT: lora_A.weight is
    [r, d_in], bound = 1/sqrt(d_in)).  B: [r, d_out] zeros (PEFT) or normal(0, 0.02) in
    test mode so that dA != 0 on step 1 (SURVEY.md §8c parity protocol)."""
    g = torch.Generator().manual_seed(seed)
    d = cfg.d_model
    out = {}
    for l in range(cfg.n_layer):
        for t in lora.targets:
            d_in, d_out = {"c_attn": (d, 3 * d), "attn_proj": (d, d),
                           "c_fc": (d, cfg.d_ff), "mlp_proj": (cfg.d_ff, d)}[t]
            bound = 1.0 / math.sqrt(d_in)
            A = (torch.rand(d_in, lora.r, generator=g, dtype=torch.float32) * 2 - 1) * bound
            if b_init == "zero":
                B = torch.zeros(lora.r, d_out)
            else:
                B = torch.randn(lora.r, d_out, generator=g, dtype=torch.float32) * 0.02
            out[f"h{l}.{t}.lora_A"] = A
            out[f"h{l}.{t}.lora_B"] = B
    return out


def synthetic_batch(cfg: GPT2Config, batch: int, seq: int, seed: int = 1234, rank: int = 0):
    """ids uniform in [0, vocab); labels = ids (HF causal-LM convention: shift inside loss)."""
    g = torch.Generator().manual_seed(seed + rank)
    ids = torch.randint(0, cfg.vocab, (batch, seq), generator=g, dtype=torch.int64)
    return ids, ids.clone()


# --------------------------------------------------------------------------------------
# the model, written out op by op
# --------------------------------------------------------------------------------------
def gelu_new(x: torch.Tensor) -> torch.Tensor:
    """GPT-2's tanh GELU ("gelu_new")."""
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * x.pow(3))))


def _linear(x, p, lp, name, lora: LoRAConfig | None):
    """Conv1D ([in,out] weight) plus optional LoRA branch."""
    y = x @ p[name + ".w"] + p[name + ".b"]
    if lora is not None and lp is not None and (name + ".lora_A") in lp:
        y = y + lora.scale * ((x @ lp[name + ".lora_A"]) @ lp[name + ".lora_B"])
    return y


def causal_attention(qkv: torch.Tensor, n_head: int) -> torch.Tensor:
    """qkv [B,T,3d] -> [B,T,d]; explicit matmul + masked softmax (no SDPA library call)."""
    B, T, d3 = qkv.shape
    d = d3 // 3
    dh = d // n_head
    q, k, v = qkv.split(d, dim=2)
    q = q.view(B, T, n_head, dh).transpose(1, 2)
    k = k.view(B, T, n_head, dh).transpose(1, 2)
    v = v.view(B, T, n_head, dh).transpose(1, 2)
    att = (q @ k.transpose(-1, -2)) / math.sqrt(dh)
    mask = torch.ones(T, T, dtype=torch.bool, device=qkv.device).tril()
    att = att.masked_fill(~mask, float("-inf"))
    att = torch.softmax(att, dim=-1)
    y = att @ v
    return y.transpose(1, 2).contiguous().view(B, T, d)


def forward(p: dict, lp: dict | None, ids: torch.Tensor, cfg: GPT2Config,
            lora: LoRAConfig | None, taps: dict | None = None) -> torch.Tensor:
    """Returns logits [B,T,V].  `taps` (optional dict) receives named intermediates."""
    B, T = ids.shape
    pos = torch.arange(T, device=ids.device)
    x = F.embedding(ids, p["wte"]) + F.embedding(pos, p["wpe"])[None]
    d = cfg.d_model
    for l in range(cfg.n_layer):
        h = F.layer_norm(x, (d,), p[f"h{l}.ln1.g"], p[f"h{l}.ln1.b"], cfg.ln_eps)
        qkv = _linear(h, p, lp, f"h{l}.c_attn", lora)
        a = causal_attention(qkv, cfg.n_head)
        x = x + _linear(a, p, lp, f"h{l}.attn_proj", lora)
        h = F.layer_norm(x, (d,), p[f"h{l}.ln2.g"], p[f"h{l}.ln2.b"], cfg.ln_eps)
        u = gelu_new(_linear(h, p, lp, f"h{l}.c_fc", lora))
        x = x + _linear(u, p, lp, f"h{l}.mlp_proj", lora)
        if taps is not None:
            taps[f"h{l}.out"] = x
    x = F.layer_norm(x, (d,), p["lnf.g"], p["lnf.b"], cfg.ln_eps)
    if taps is not None:
        taps["lnf.out"] = x
    return x @ p["wte"].t()  # tied lm_head


def causal_lm_loss(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """HF convention: predict token t+1 from position t; mean over non-ignored targets."""
    V = logits.shape[-1]
    return F.cross_entropy(logits[:, :-1].reshape(-1, V).float(), labels[:, 1:].reshape(-1),
                           ignore_index=ignore_index)


# --------------------------------------------------------------------------------------
# kernel-level restatements (what each fused CUDA kernel must reproduce)
# --------------------------------------------------------------------------------------
def softmax_xent_ref(z: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100):
    """z [M,V] fp32, targets [M] int64.  Returns (mean loss, dz) with
    dz = (softmax(z) - onehot) / N_valid on valid rows, 0 on ignored rows."""
    z = z.float()
    valid = targets != ignore_index
    n = int(valid.sum().item())
    lse = torch.logsumexp(z, dim=-1)
    tgt = targets.clamp(min=0)
    picked = z.gather(1, tgt[:, None])[:, 0]
    loss_rows = torch.where(valid, lse - picked, torch.zeros_like(lse))
    loss = loss_rows.sum() / max(n, 1)
    dz = torch.softmax(z, dim=-1)
    dz.scatter_add_(1, tgt[:, None], -torch.ones_like(picked)[:, None])
    dz = dz * valid[:, None] / max(n, 1)
    return loss, dz


def adamw_ref(p, g, m, v, step: int, o: AdamWConfig, grad_scale: float = 1.0):
    """One decoupled-weight-decay Adam update, the formula of torch.optim.AdamW
    (non-fused, non-amsgrad), written out.  Returns new (p, m, v)."""
    g = g.float() * grad_scale
    p = p * (1.0 - o.lr * o.weight_decay)
    m = o.beta1 * m + (1.0 - o.beta1) * g
    v = o.beta2 * v + (1.0 - o.beta2) * g * g
    bc1 = 1.0 - o.beta1 ** step
    bc2 = 1.0 - o.beta2 ** step
    denom = v.sqrt() / math.sqrt(bc2) + o.eps
    p = p - (o.lr / bc1) * (m / denom)
    return p, m, v


# --------------------------------------------------------------------------------------
# the step
# --------------------------------------------------------------------------------------
class OracleTrainer:
    """fwd + shifted CE + autograd bwd + torch.optim.AdamW(foreach=False, fused=False)."""

    def __init__(self, cfg: GPT2Config, lora: LoRAConfig | None, opt: AdamWConfig,
                 base: dict, adapters: dict | None, mode: str = "lora"):
        assert mode in ("lora", "full")
        self.cfg, self.lora, self.opt_cfg, self.mode = cfg, lora, opt, mode
        self.base = {k: v.clone().float() for k, v in base.items()}
        self.adapters = {k: v.clone().float() for k, v in (adapters or {}).items()}
        if mode == "lora":
            self.trainable = self.adapters
        else:
            self.trainable = self.base
            self.adapters, self.lora = {}, None
        for t in self.trainable.values():
            t.requires_grad_(True)
        self.optim = torch.optim.AdamW(list(self.trainable.values()), lr=opt.lr,
                                       betas=(opt.beta1, opt.beta2), eps=opt.eps,
                                       weight_decay=opt.weight_decay, foreach=False, fused=False)
        self.last_grads: dict = {}

    def loss_only(self, ids, labels) -> float:
        with torch.no_grad():
            return float(causal_lm_loss(forward(self.base, self.adapters, ids, self.cfg, self.lora), labels))

    def step(self, ids, labels, world_grads: list | None = None) -> float:
        """One optimizer step.  `world_grads`: optional list of per-rank (ids, labels) to
        emulate data-parallel averaging (mean of per-rank mean losses' gradients)."""
        self.optim.zero_grad(set_to_none=True)
        batches = world_grads if world_grads is not None else [(ids, labels)]
        total = 0.0
        for bi, bl in batches:
            logits = forward(self.base, self.adapters, bi, self.cfg, self.lora)
            loss = causal_lm_loss(logits, bl) / len(batches)
            loss.backward()
            total += float(loss.detach())
        if self.opt_cfg.max_grad_norm > 0:
            torch.nn.utils.clip_grad_norm_(list(self.trainable.values()), self.opt_cfg.max_grad_norm)
        self.last_grads = {k: t.grad.detach().clone() for k, t in self.trainable.items()}
        self.optim.step()
        return total


# --------------------------------------------------------------------------------------
# pinning the restatement against an independent implementation
# --------------------------------------------------------------------------------------
def hf_model(cfg: GPT2Config, p: dict):
    """transformers' GPT2LMHeadModel built from config (no pretrained weights needed) holding this module's
    parameters `p`: the independent implementation the restatement is pinned against."""
    from transformers import GPT2Config as HFConfig, GPT2LMHeadModel

    hf = GPT2LMHeadModel(HFConfig(
        vocab_size=cfg.vocab, n_positions=cfg.ctx, n_embd=cfg.d_model, n_layer=cfg.n_layer,
        n_head=cfg.n_head, n_inner=cfg.d_ff, activation_function="gelu_new",
        resid_pdrop=0.0, embd_pdrop=0.0, attn_pdrop=0.0, layer_norm_epsilon=cfg.ln_eps,
        attn_implementation="eager")).float().eval()
    sd = hf.state_dict()
    with torch.no_grad():
        sd["transformer.wte.weight"].copy_(p["wte"])
        sd["transformer.wpe.weight"].copy_(p["wpe"])
        for l in range(cfg.n_layer):
            pre = f"transformer.h.{l}."
            for mine, theirs in (("ln1", "ln_1"), ("ln2", "ln_2")):
                sd[pre + theirs + ".weight"].copy_(p[f"h{l}.{mine}.g"])
                sd[pre + theirs + ".bias"].copy_(p[f"h{l}.{mine}.b"])
            for mine, theirs in (("c_attn", "attn.c_attn"), ("attn_proj", "attn.c_proj"),
                                 ("c_fc", "mlp.c_fc"), ("mlp_proj", "mlp.c_proj")):
                sd[pre + theirs + ".weight"].copy_(p[f"h{l}.{mine}.w"])
                sd[pre + theirs + ".bias"].copy_(p[f"h{l}.{mine}.b"])
        sd["transformer.ln_f.weight"].copy_(p["lnf.g"])
        sd["transformer.ln_f.bias"].copy_(p["lnf.b"])
    hf.tie_weights()
    return hf


def hf_crosscheck(cfg: GPT2Config = GPT2_TINY, batch: int = 2, seq: int = 32, seed: int = 0) -> dict:
    """Compare logits / loss / gradients of this restatement with transformers' GPT-2 on the same weights."""
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
        "wte_grad_max_abs_diff": float((mine["wte"].grad - hf.transformer.wte.weight.grad).abs().max()),
        "c_attn0_grad_max_abs_diff": float(
            (mine["h0.c_attn.w"].grad - hf.transformer.h[0].attn.c_attn.weight.grad).abs().max()),
    }


if __name__ == "__main__":  # pragma: no cover
    print(hf_crosscheck())
