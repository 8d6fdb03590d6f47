"""The oracle itself, on CPU: pinned against transformers' GPT-2 (an independent
implementation), against torch's own AdamW / cross_entropy, and against the committed golden
fixture (so that a later edit of the oracle cannot silently move the goalposts)."""
import math
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import gpt2_lora_ref as ref

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "gpt2_tiny_lora_step.pt")


def test_restatement_matches_hf_gpt2():
    r = ref.hf_crosscheck(ref.GPT2_TINY, batch=2, seq=32)
    assert r["logits_max_abs_diff"] < 1e-5
    assert abs(r["loss_mine"] - r["loss_hf"]) < 1e-5
    assert r["wte_grad_max_abs_diff"] < 1e-6 and r["c_attn0_grad_max_abs_diff"] < 1e-6


def test_restatement_matches_hf_gpt2_real_head_dim():
    cfg = ref.GPT2Config(n_layer=1, d_model=128, n_head=2, d_ff=512, vocab=1000, ctx=64)  # dh = 64 as in GPT-2
    r = ref.hf_crosscheck(cfg, batch=1, seq=48)
    assert r["logits_max_abs_diff"] < 1e-5 and abs(r["loss_mine"] - r["loss_hf"]) < 1e-5


def test_initial_loss_is_ln_vocab():
    cfg = ref.GPT2_TINY
    p = ref.init_base_params(cfg, 0)
    ids, labels = ref.synthetic_batch(cfg, 4, 32)
    loss = float(ref.causal_lm_loss(ref.forward(p, None, ids, cfg, None), labels))
    assert abs(loss - math.log(cfg.vocab)) < 0.1


def test_lora_branch_equals_merged_weights():
    cfg, lora = ref.GPT2_TINY, ref.LoRAConfig(r=4, alpha=8.0)
    p = ref.init_base_params(cfg, 0, randomize_all=True)
    ad = ref.init_lora_params(cfg, lora, 1, b_init="normal")
    ids, _ = ref.synthetic_batch(cfg, 2, 16)
    y = ref.forward(p, ad, ids, cfg, lora)
    merged = dict(p)
    for l in range(cfg.n_layer):
        merged[f"h{l}.c_attn.w"] = p[f"h{l}.c_attn.w"] + lora.scale * ad[f"h{l}.c_attn.lora_A"] @ ad[f"h{l}.c_attn.lora_B"]
    torch.testing.assert_close(y, ref.forward(merged, None, ids, cfg, None), rtol=1e-4, atol=1e-5)


def test_peft_init_zero_b_means_no_change_and_zero_dA():
    cfg, lora = ref.GPT2_TINY, ref.LoRAConfig()
    p = ref.init_base_params(cfg, 0)
    ad = ref.init_lora_params(cfg, lora, 1)
    ids, labels = ref.synthetic_batch(cfg, 2, 16)
    torch.testing.assert_close(ref.forward(p, ad, ids, cfg, lora), ref.forward(p, None, ids, cfg, None))
    tr = ref.OracleTrainer(cfg, lora, ref.AdamWConfig(), p, ad)
    tr.step(ids, labels)
    assert float(tr.last_grads["h0.c_attn.lora_A"].abs().max()) == 0.0
    assert float(tr.last_grads["h0.c_attn.lora_B"].abs().max()) > 0.0
    bound = 1 / math.sqrt(cfg.d_model)
    assert float(ad["h0.c_attn.lora_A"].abs().max()) <= bound


def test_softmax_xent_ref_matches_torch():
    g = torch.Generator().manual_seed(0)
    z = torch.randn(9, 101, generator=g, requires_grad=True)
    t = torch.randint(0, 101, (9,), generator=g)
    t[3] = -100
    loss, dz = ref.softmax_xent_ref(z.detach(), t)
    lt = F.cross_entropy(z, t, ignore_index=-100)
    lt.backward()
    assert abs(float(loss) - float(lt)) < 1e-6
    torch.testing.assert_close(dz, z.grad, rtol=1e-5, atol=1e-7)
    loss0, dz0 = ref.softmax_xent_ref(z.detach(), torch.full((9,), -100))
    assert float(loss0) == 0.0 and float(dz0.abs().max()) == 0.0


def test_adamw_ref_matches_torch_optim():
    g = torch.Generator().manual_seed(1)
    o = ref.AdamWConfig(lr=1e-2, beta1=0.8, beta2=0.9, weight_decay=0.05)
    p = torch.randn(257, generator=g)
    pt = p.clone().requires_grad_(True)
    opt = torch.optim.AdamW([pt], lr=o.lr, betas=(o.beta1, o.beta2), eps=o.eps, weight_decay=o.weight_decay,
                            foreach=False, fused=False)
    m, v = torch.zeros(257), torch.zeros(257)
    for step in range(1, 5):
        grad = torch.randn(257, generator=g)
        pt.grad = grad.clone()
        opt.step()
        p, m, v = ref.adamw_ref(p, grad, m, v, step, o)
        torch.testing.assert_close(p, pt.detach(), rtol=1e-6, atol=1e-7)


def test_data_parallel_emulation_is_mean_of_rank_grads():
    cfg, lora = ref.GPT2_TINY, ref.LoRAConfig()
    p = ref.init_base_params(cfg, 0)
    ad = ref.init_lora_params(cfg, lora, 1, b_init="normal")
    b0, b1 = ref.synthetic_batch(cfg, 2, 16, rank=0), ref.synthetic_batch(cfg, 2, 16, rank=1)
    grads = []
    for b in (b0, b1):
        tr = ref.OracleTrainer(cfg, lora, ref.AdamWConfig(), p, ad)
        tr.step(*b)
        grads.append(tr.last_grads)
    tr = ref.OracleTrainer(cfg, lora, ref.AdamWConfig(), p, ad)
    tr.step(None, None, world_grads=[b0, b1])
    for k in grads[0]:
        torch.testing.assert_close(tr.last_grads[k], 0.5 * (grads[0][k] + grads[1][k]), rtol=1e-4, atol=1e-8)


def test_golden_fixture_reproduces():
    gold = torch.load(GOLDEN)
    cfg, lora, opt = ref.GPT2_TINY, ref.LoRAConfig(r=8, alpha=16.0), ref.AdamWConfig(lr=1e-3)
    tr = ref.OracleTrainer(cfg, lora, opt, ref.init_base_params(cfg, 0, randomize_all=True),
                           ref.init_lora_params(cfg, lora, 1, b_init="normal"))
    for s, l_gold in enumerate(gold["losses"]):
        ids, labels = ref.synthetic_batch(cfg, 2, 32, seed=1234 + s)
        assert torch.equal(ids, gold["ids"][s])
        assert abs(tr.step(ids, labels) - l_gold) < 1e-5
        if s == 0:
            for k, g in gold["grads_step1"].items():
                torch.testing.assert_close(tr.last_grads[k], g, rtol=1e-4, atol=1e-7)
    for k, t in gold["adapters_after"].items():
        torch.testing.assert_close(tr.trainable[k].detach(), t, rtol=1e-4, atol=1e-6)
    x = gold["xent"]
    loss, dz = ref.softmax_xent_ref(x["z"], x["targets"])
    assert abs(float(loss) - x["loss"]) < 1e-6
    torch.testing.assert_close(dz, x["dz"])
    a = gold["adamw"]
    p1, m1, v1 = ref.adamw_ref(a["p0"], a["g0"], torch.zeros(64), torch.zeros(64), 1, opt, grad_scale=a["grad_scale"])
    torch.testing.assert_close(p1, a["p1"])
