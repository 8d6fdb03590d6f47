"""Parameter construction for the payload: pretrained base weights from the dataset mount when the job
names them (weights.load_base_weights), random init otherwise (benchmarks, smoke runs).

Naming/layout is HF GPT-2's: Conv1D weights are [in, out]; see GPT2LoRAStep._upload_base.
"""
from __future__ import annotations

import math

import torch

from .gpt2 import GPT2Dims, LoRASpec
from .llama import LLAMA_3_2_1B, LLAMA_3_8B, LlamaDims

MODELS = {
    "gpt2-124m": GPT2Dims(12, 768, 12, 3072, 50257, 1024, 1e-5),
    "gpt2-tiny": GPT2Dims(2, 64, 2, 256, 503, 64, 1e-5),
    "llama-3.2-1b": LLAMA_3_2_1B,
    "llama-3-8b": LLAMA_3_8B,
    "llama-tiny": LlamaDims(n_layer=2, d_model=64, n_head=4, n_kv_head=2, d_ff=128, vocab=503, ctx=128,
                            rope_theta=10000.0, rope_llama3=False, tie_embeddings=True),
}


def is_llama(dims) -> bool:
    return isinstance(dims, LlamaDims)


def init_gpt2(dims: GPT2Dims, seed: int, device="cpu") -> dict:
    g = torch.Generator(device=device).manual_seed(seed)
    d, ff = dims.d_model, dims.d_ff

    def w(*shape):
        return torch.empty(*shape, device=device).normal_(0.0, 0.02, generator=g)

    p = {"wte": w(dims.vocab, d), "wpe": w(dims.ctx, d),
         "lnf.g": torch.ones(d, device=device), "lnf.b": torch.zeros(d, device=device)}
    for l in range(dims.n_layer):
        for name, (i, o) in (("c_attn", (d, 3 * d)), ("attn_proj", (d, d)), ("c_fc", (d, ff)),
                             ("mlp_proj", (ff, d))):
            p[f"h{l}.{name}.w"] = w(i, o)
            p[f"h{l}.{name}.b"] = torch.zeros(o, device=device)
        for ln in ("ln1", "ln2"):
            p[f"h{l}.{ln}.g"] = torch.ones(d, device=device)
            p[f"h{l}.{ln}.b"] = torch.zeros(d, device=device)
    return p


def init_lora(dims: GPT2Dims, lora: LoRASpec, seed: int, device="cpu") -> dict:
    """PEFT convention: A uniform(-1/sqrt(d_in), 1/sqrt(d_in)), B zero."""
    g = torch.Generator(device=device).manual_seed(seed)
    d = dims.d_model
    bound = 1.0 / math.sqrt(d)
    out = {}
    for l in range(dims.n_layer):
        out[f"h{l}.c_attn.lora_A"] = torch.empty(d, lora.r, device=device).uniform_(-bound, bound, generator=g)
        out[f"h{l}.c_attn.lora_B"] = torch.zeros(lora.r, 3 * d, device=device)
    return out


def init_llama(dims: LlamaDims, seed: int, device="cpu") -> dict:
    """nn.Linear layout [out, in]; normal(0, 0.02) weights, unit RMSNorm gains."""
    g = torch.Generator(device=device).manual_seed(seed)
    d, ff, dh = dims.d_model, dims.d_ff, dims.d_head

    def w(*shape):
        return torch.empty(*shape, device=device).normal_(0.0, 0.02, generator=g)

    p = {"embed": w(dims.vocab, d), "norm.g": torch.ones(d, device=device)}
    for l in range(dims.n_layer):
        p[f"l{l}.q_proj.w"] = w(dims.n_head * dh, d)
        p[f"l{l}.k_proj.w"] = w(dims.n_kv_head * dh, d)
        p[f"l{l}.v_proj.w"] = w(dims.n_kv_head * dh, d)
        p[f"l{l}.o_proj.w"] = w(d, dims.n_head * dh)
        p[f"l{l}.gate_proj.w"] = w(ff, d)
        p[f"l{l}.up_proj.w"] = w(ff, d)
        p[f"l{l}.down_proj.w"] = w(d, ff)
        p[f"l{l}.attn_norm.g"] = torch.ones(d, device=device)
        p[f"l{l}.mlp_norm.g"] = torch.ones(d, device=device)
    if not dims.tie_embeddings:
        p["lm_head.w"] = w(dims.vocab, d)
    return p


def init_lora_llama(dims: LlamaDims, lora: LoRASpec, seed: int, device="cpu") -> dict:
    """PEFT convention on q_proj and v_proj: lora_A [r, in] uniform(+-1/sqrt(in)), lora_B [out, r] zero."""
    g = torch.Generator(device=device).manual_seed(seed)
    d, dh = dims.d_model, dims.d_head
    bound = 1.0 / math.sqrt(d)
    out = {}
    for l in range(dims.n_layer):
        for t, n_out in (("q_proj", dims.n_head * dh), ("v_proj", dims.n_kv_head * dh)):
            out[f"l{l}.{t}.lora_A"] = torch.empty(lora.r, d, device=device).uniform_(-bound, bound, generator=g)
            out[f"l{l}.{t}.lora_B"] = torch.zeros(n_out, lora.r, device=device)
    return out


def build_engine(model: str, lora: LoRASpec, opt, *, batch: int, seq: int, device, seed: int = 0, world_size: int = 1,
                 grad_sync=None, randomize_b: bool = False, mode: str = "lora", attention: str = "auto",
                 overlap_optimizer: bool = False, base_weights: str = ""):
    """Model (+ adapters) of the named architecture on `device`, wrapped in its step engine.  The base comes
    from `base_weights` (a Hugging Face checkpoint file / directory, see weights.py) or, if empty, is drawn at
    random.  mode "lora" trains adapters on a frozen base; mode "full" (Llama family) trains every tensor."""
    from . import gpt2, llama
    dims = MODELS[model]
    loaded = None
    if base_weights:
        from . import weights
        loaded = weights.load_base_weights(base_weights, dims, is_llama(dims))
    if mode == "full":
        if not is_llama(dims):
            raise ValueError("mode=full is implemented for the Llama family (BASELINE.json configs[4])")
        from . import llama_full
        return llama_full.LlamaFullStep(dims, opt, loaded, batch=batch, seq=seq, device=device, world_size=world_size,
                                        grad_sync=grad_sync, seed=seed, attention=attention,
                                        overlap_optimizer=overlap_optimizer), dims
    if mode != "lora":
        raise ValueError(f"unknown mode {mode!r}")
    if is_llama(dims):
        base, ad = loaded or init_llama(dims, seed, device), init_lora_llama(dims, lora, seed + 1, device)
        cls = llama.LlamaLoRAStep
    else:
        base, ad = loaded or init_gpt2(dims, seed, device), init_lora(dims, lora, seed + 1, device)
        cls = gpt2.GPT2LoRAStep
    if randomize_b:  # benchmarks: every gradient path carries non-trivial data from step 1
        gen = torch.Generator(device=device).manual_seed(seed + 99)
        for k in ad:
            if k.endswith("lora_B"):
                ad[k].normal_(0.0, 0.02, generator=gen)
    return cls(dims, lora, opt, base, ad, batch=batch, seq=seq, device=device, world_size=world_size,
               grad_sync=grad_sync, attention=attention), dims
