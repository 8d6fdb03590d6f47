"""The Llama oracle on CPU: pinned against transformers' LlamaForCausalLM (independent
implementation) with and without the llama3 RoPE rescaling and with grouped-query attention."""
import torch

from oracle import llama_lora_ref as ref


def test_llama_restatement_matches_hf():
    r = ref.hf_crosscheck(ref.LLAMA_TINY, batch=2, seq=24)
    assert r["logits_max_abs_diff"] < 1e-5 and abs(r["loss_mine"] - r["loss_hf"]) < 1e-5
    assert r["q0_grad_max_abs_diff"] < 1e-6 and r["down0_grad_max_abs_diff"] < 1e-6


def test_llama3_rope_scaling_and_untied_head_match_hf():
    cfg = ref.LlamaConfig(n_layer=1, d_model=128, n_head=4, n_kv_head=2, d_ff=256, vocab=1000, ctx=16384,
                          rope_llama3=True, rope_original_ctx=64, tie_embeddings=False)
    r = ref.hf_crosscheck(cfg, batch=1, seq=96)   # positions beyond the "original" context: scaling active
    assert r["logits_max_abs_diff"] < 1e-5 and abs(r["loss_mine"] - r["loss_hf"]) < 1e-5
    inv_plain = ref.rope_inv_freq(ref.LlamaConfig(d_model=128, n_head=4, rope_llama3=False))
    assert not torch.allclose(ref.rope_inv_freq(cfg), inv_plain)


def test_lora_on_q_and_v_equals_merged_weights_and_zero_b_is_identity():
    cfg, lora = ref.LLAMA_TINY, ref.LoRAConfig(r=4, alpha=8.0)
    p = ref.init_base_params(cfg, 0, randomize_all=True)
    ids, labels = ref.synthetic_batch(cfg, 2, 16)
    ad0 = ref.init_lora_params(cfg, lora, 1)                     # PEFT init: B = 0
    torch.testing.assert_close(ref.forward(p, ad0, ids, cfg, lora), ref.forward(p, None, ids, cfg, None))
    ad = ref.init_lora_params(cfg, lora, 1, b_init="normal")
    merged = dict(p)
    for l in range(cfg.n_layer):
        for t in lora.targets:
            merged[f"l{l}.{t}.w"] = p[f"l{l}.{t}.w"] + lora.scale * ad[f"l{l}.{t}.lora_B"] @ ad[f"l{l}.{t}.lora_A"]
    torch.testing.assert_close(ref.forward(p, ad, ids, cfg, lora), ref.forward(merged, None, ids, cfg, None),
                               rtol=1e-4, atol=1e-5)
    tr = ref.OracleTrainer(cfg, lora, ref.AdamWConfig(), p, ad0)
    tr.step(ids, labels)
    assert float(tr.last_grads["l0.q_proj.lora_A"].abs().max()) == 0.0
    assert float(tr.last_grads["l0.v_proj.lora_B"].abs().max()) > 0.0
