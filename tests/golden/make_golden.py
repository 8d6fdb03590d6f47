"""Generates tests/golden/gpt2_tiny_lora_step.pt from the CPU oracle (run in the build
container: `python tests/golden/make_golden.py`).  The reference repository holds no golden
vectors for the training step, so these are the oracle's outputs — itself
pinned against transformers' GPT2LMHeadModel by tests/test_oracle.py — frozen to a file."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import gpt2_lora_ref as ref  # noqa: E402


def main():
    cfg = ref.GPT2_TINY
    lora = ref.LoRAConfig(r=8, alpha=16.0)
    opt = ref.AdamWConfig(lr=1e-3)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = ref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    tr = ref.OracleTrainer(cfg, lora, opt, base, ad)
    losses, ids_all, grads1 = [], [], None
    for s in range(3):
        ids, labels = ref.synthetic_batch(cfg, 2, 32, seed=1234 + s)
        ids_all.append(ids)
        losses.append(tr.step(ids, labels))
        if s == 0:
            grads1 = {k: v.clone() for k, v in tr.last_grads.items()}
    # kernel-level vectors
    g = torch.Generator().manual_seed(7)
    z = (torch.randn(6, 503, generator=g) * 3).to(torch.bfloat16).float()
    tg = torch.randint(0, 503, (6,), generator=g)
    tg[2] = -100
    xl, xdz = ref.softmax_xent_ref(z, tg)
    p0, g0 = torch.randn(64, generator=g), torch.randn(64, generator=g)
    p1, m1, v1 = ref.adamw_ref(p0, g0, torch.zeros(64), torch.zeros(64), 1, opt, grad_scale=0.5)
    out = {
        "losses": losses, "ids": ids_all, "grads_step1": grads1,
        "adapters_after": {k: v.detach().clone() for k, v in tr.trainable.items()},
        "xent": {"z": z, "targets": tg, "loss": float(xl), "dz": xdz},
        "adamw": {"p0": p0, "g0": g0, "p1": p1, "m1": m1, "v1": v1, "grad_scale": 0.5},
        "torch_version": str(torch.__version__),
    }
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "gpt2_tiny_lora_step.pt")
    torch.save(out, path)
    print("wrote", path, "losses", losses)


if __name__ == "__main__":
    main()
