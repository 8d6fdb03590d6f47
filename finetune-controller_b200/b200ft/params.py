"""Parameter construction for the payload (random init — the image has no network, so no
pretrained checkpoint can be fetched; a checkpoint under the dataset mount is used if present).

Naming/layout is HF GPT-2's: Conv1D weights are [in, out]; see GPT2LoRAStep._upload_base.
"""
from __future__ import annotations

import math

import torch

from .gpt2 import GPT2Dims, LoRASpec

MODELS = {
    "gpt2-124m": GPT2Dims(12, 768, 12, 3072, 50257, 1024, 1e-5),
    "gpt2-tiny": GPT2Dims(2, 64, 2, 256, 503, 64, 1e-5),
}


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
