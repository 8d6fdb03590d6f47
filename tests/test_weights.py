"""Weight files either side of the step, on CPU:

  * a checkpoint in Hugging Face naming (what `transformers` itself writes) loads into the engines' naming;
  * the adapters the payload writes as adapter_model.safetensors, merged the way PEFT merges them
    (delta = lora_B.weight @ lora_A.weight * alpha/r, transposed for a fan_in_fan_out Conv1D target) into
    transformers' own model, give the logits of the oracle's forward with the adapters attached;
  * the oracle's LoRA GRADIENTS are pinned independently: transformers' model with merged weights gives
    dL/dW', and the chain rule dA = s (dL/dW') B^T, dB = s A^T (dL/dW') must equal the oracle's autograd."""
import json
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
from b200ft import params, weights  # noqa: E402
from oracle import gpt2_lora_ref as gref  # noqa: E402
from oracle import llama_lora_ref as lref  # noqa: E402

safetensors = pytest.importorskip("safetensors")
from safetensors.torch import load_file, save_file  # noqa: E402

GPT_DIMS, LLAMA_DIMS = params.MODELS["gpt2-tiny"], params.MODELS["llama-tiny"]
LLAMA_CFG = lref.LlamaConfig(n_layer=2, d_model=64, n_head=4, n_kv_head=2, d_ff=128, vocab=503, ctx=128,
                             rope_theta=10000.0, rope_llama3=False, tie_embeddings=True)


def _hf_state(hf, drop_tied):
    sd = {k: v.detach().clone().contiguous() for k, v in hf.state_dict().items()}
    if drop_tied:
        sd.pop("lm_head.weight", None)          # safetensors refuses shared storage; HF drops the tied head too
    return sd


def test_hf_gpt2_checkpoint_loads_into_engine_naming(tmp_path):
    p = gref.init_base_params(gref.GPT2_TINY, 3, randomize_all=True)
    sd = _hf_state(gref.hf_model(gref.GPT2_TINY, p), drop_tied=True)
    assert "transformer.h.0.attn.c_attn.weight" in sd
    save_file(sd, str(tmp_path / "model.safetensors"))
    for path in (tmp_path / "model.safetensors", tmp_path):                       # a file, or the directory holding it
        got = weights.load_base_weights(str(path), GPT_DIMS, llama=False)
        assert set(got) == set(p)
        for k in p:
            assert torch.equal(got[k], p[k]), k
    # a torch.save state dict without the "transformer." prefix (GPT2Model) is accepted too
    torch.save({k.replace("transformer.", ""): v for k, v in sd.items()}, tmp_path / "m.pt")
    got = weights.load_base_weights(str(tmp_path / "m.pt"), GPT_DIMS, llama=False)
    assert torch.equal(got["h1.mlp_proj.w"], p["h1.mlp_proj.w"])


def test_hf_llama_checkpoint_loads_sharded_and_untied(tmp_path):
    p = lref.init_base_params(LLAMA_CFG, 4, randomize_all=True)
    sd = _hf_state(lref.hf_model(LLAMA_CFG, p), drop_tied=True)
    keys = sorted(sd)
    save_file({k: sd[k] for k in keys[: len(keys) // 2]}, str(tmp_path / "model-00001-of-00002.safetensors"))
    save_file({k: sd[k] for k in keys[len(keys) // 2:]}, str(tmp_path / "model-00002-of-00002.safetensors"))
    got = weights.load_base_weights(str(tmp_path), LLAMA_DIMS, llama=True)
    assert set(got) == set(p)
    for k in p:
        assert torch.equal(got[k], p[k]), k
    import dataclasses
    cfg_u = dataclasses.replace(LLAMA_CFG, tie_embeddings=False)
    dims_u = dataclasses.replace(LLAMA_DIMS, tie_embeddings=False)
    pu = lref.init_base_params(cfg_u, 5, randomize_all=True)
    d2 = tmp_path / "untied"; d2.mkdir()
    save_file(_hf_state(lref.hf_model(cfg_u, pu), drop_tied=False), str(d2 / "model.safetensors"))
    got = weights.load_base_weights(str(d2), dims_u, llama=True)
    assert torch.equal(got["lm_head.w"], pu["lm_head.w"]) and not torch.equal(got["lm_head.w"], got["embed"])


def test_wrong_checkpoint_is_refused(tmp_path):
    p = gref.init_base_params(gref.GPT2_TINY, 3)
    save_file(_hf_state(gref.hf_model(gref.GPT2_TINY, p), drop_tied=True), str(tmp_path / "model.safetensors"))
    with pytest.raises(KeyError, match="Llama"):
        weights.load_base_weights(str(tmp_path), LLAMA_DIMS, llama=True)          # GPT-2 file, Llama job
    with pytest.raises(KeyError, match="h.2."):
        weights.load_base_weights(str(tmp_path), params.MODELS["gpt2-124m"], llama=False)   # 2-layer file, 12-layer job
    import dataclasses
    with pytest.raises(ValueError, match="shape"):
        weights.load_base_weights(str(tmp_path), dataclasses.replace(GPT_DIMS, d_ff=512), llama=False)
    with pytest.raises(FileNotFoundError):
        weights.load_base_weights(str(tmp_path / "empty_dir_that_is_a_file.safetensors"), GPT_DIMS, llama=False)


def _merge_peft_into_hf(hf, tensors: dict, scale: float, fan_in_fan_out: bool) -> None:
    """What PEFT's merge does, written out: module.weight += T(lora_B.weight @ lora_A.weight) * scale."""
    sd = hf.state_dict()
    mods = sorted({k[len("base_model.model."):].rsplit(".lora_", 1)[0] for k in tensors})
    with torch.no_grad():
        for m in mods:
            a, b = tensors[f"base_model.model.{m}.lora_A.weight"], tensors[f"base_model.model.{m}.lora_B.weight"]
            delta = (b @ a) * scale
            sd[m + ".weight"].add_(delta.t() if fan_in_fan_out else delta)


def test_peft_adapter_file_merged_into_hf_gpt2_gives_the_oracles_logits(tmp_path):
    cfg, lora = gref.GPT2_TINY, gref.LoRAConfig(r=4, alpha=12.0)
    p = gref.init_base_params(cfg, 0, randomize_all=True)
    ad = gref.init_lora_params(cfg, lora, 1, b_init="normal")
    weights.save_peft_adapter(str(tmp_path), ad, model="gpt2-tiny", llama=False, r=lora.r, alpha=lora.alpha, steps=7)
    t = load_file(str(tmp_path / "adapter_model.safetensors"))
    c = json.load(open(tmp_path / "adapter_config.json"))
    assert c["peft_type"] == "LORA" and c["target_modules"] == ["c_attn"] and c["fan_in_fan_out"] is True
    assert c["r"] == 4 and c["lora_alpha"] == 12.0 and c["task_type"] == "CAUSAL_LM"
    a0 = t["base_model.model.transformer.h.0.attn.c_attn.lora_A.weight"]
    assert a0.shape == (4, cfg.d_model) and t["base_model.model.transformer.h.1.attn.c_attn.lora_B.weight"].shape == (3 * cfg.d_model, 4)
    hf = gref.hf_model(cfg, p)
    _merge_peft_into_hf(hf, t, c["lora_alpha"] / c["r"], c["fan_in_fan_out"])
    ids, _ = gref.synthetic_batch(cfg, 2, 24)
    with torch.no_grad():
        want = gref.forward(p, ad, ids, cfg, lora)
        got = hf(input_ids=ids).logits
    torch.testing.assert_close(got, want, rtol=1e-4, atol=2e-5)


def test_peft_adapter_file_merged_into_hf_llama_gives_the_oracles_logits(tmp_path):
    cfg, lora = LLAMA_CFG, lref.LoRAConfig(r=8, alpha=16.0)
    p = lref.init_base_params(cfg, 0, randomize_all=True)
    ad = lref.init_lora_params(cfg, lora, 1, b_init="normal")
    weights.save_peft_adapter(str(tmp_path), ad, model="llama-tiny", llama=True, r=lora.r, alpha=lora.alpha, steps=1)
    t = load_file(str(tmp_path / "adapter_model.safetensors"))
    c = json.load(open(tmp_path / "adapter_config.json"))
    assert c["target_modules"] == ["q_proj", "v_proj"] and c["fan_in_fan_out"] is False
    assert t["base_model.model.model.layers.1.self_attn.v_proj.lora_B.weight"].shape == (cfg.n_kv_head * cfg.d_head, 8)
    hf = lref.hf_model(cfg, p)
    _merge_peft_into_hf(hf, t, c["lora_alpha"] / c["r"], False)
    ids, _ = lref.synthetic_batch(cfg, 2, 24)
    with torch.no_grad():
        want = lref.forward(p, ad, ids, cfg, lora)
        got = hf(input_ids=ids).logits
    torch.testing.assert_close(got, want, rtol=1e-4, atol=2e-5)


def test_oracle_lora_gradients_pinned_through_hf_merged_weights():
    """dA, dB of the oracle (autograd through its own LoRA branch) against the chain rule applied to the
    weight gradient of an INDEPENDENT model (transformers GPT-2 holding W' = W + s A B)."""
    cfg, lora = gref.GPT2_TINY, gref.LoRAConfig(r=4, alpha=8.0)
    s = lora.scale
    p = gref.init_base_params(cfg, 0, randomize_all=True)
    ad = gref.init_lora_params(cfg, lora, 1, b_init="normal")
    ids, labels = gref.synthetic_batch(cfg, 2, 24)
    tr = gref.OracleTrainer(cfg, lora, gref.AdamWConfig(), p, ad)
    tr.step(ids, labels)
    merged = dict(p)
    for l in range(cfg.n_layer):
        merged[f"h{l}.c_attn.w"] = p[f"h{l}.c_attn.w"] + s * ad[f"h{l}.c_attn.lora_A"] @ ad[f"h{l}.c_attn.lora_B"]
    hf = gref.hf_model(cfg, merged)
    hf(input_ids=ids, labels=labels).loss.backward()
    for l in range(cfg.n_layer):
        gw = hf.transformer.h[l].attn.c_attn.weight.grad                       # Conv1D: [in, out]
        A, B = ad[f"h{l}.c_attn.lora_A"], ad[f"h{l}.c_attn.lora_B"]           # [in, r], [r, out]
        torch.testing.assert_close(tr.last_grads[f"h{l}.c_attn.lora_A"], s * gw @ B.t(), rtol=1e-4, atol=1e-7)
        torch.testing.assert_close(tr.last_grads[f"h{l}.c_attn.lora_B"], s * A.t() @ gw, rtol=1e-4, atol=1e-7)


def test_oracle_llama_lora_gradients_pinned_through_hf_merged_weights():
    cfg, lora = LLAMA_CFG, lref.LoRAConfig(r=8, alpha=16.0)
    s = lora.scale
    p = lref.init_base_params(cfg, 0, randomize_all=True)
    ad = lref.init_lora_params(cfg, lora, 1, b_init="normal")
    ids, labels = lref.synthetic_batch(cfg, 2, 24)
    tr = lref.OracleTrainer(cfg, lora, gref.AdamWConfig(), p, ad)
    tr.step(ids, labels)
    merged = dict(p)
    for l in range(cfg.n_layer):
        for t in ("q_proj", "v_proj"):     # nn.Linear [out, in]; lora_B [out, r], lora_A [r, in]
            merged[f"l{l}.{t}.w"] = p[f"l{l}.{t}.w"] + s * ad[f"l{l}.{t}.lora_B"] @ ad[f"l{l}.{t}.lora_A"]
    hf = lref.hf_model(cfg, merged)
    hf(input_ids=ids, labels=labels).loss.backward()
    for l in range(cfg.n_layer):
        for t in ("q_proj", "v_proj"):
            gw = getattr(hf.model.layers[l].self_attn, t).weight.grad            # [out, in]
            A, B = ad[f"l{l}.{t}.lora_A"], ad[f"l{l}.{t}.lora_B"]
            torch.testing.assert_close(tr.last_grads[f"l{l}.{t}.lora_A"], s * B.t() @ gw, rtol=1e-4, atol=1e-7)
            torch.testing.assert_close(tr.last_grads[f"l{l}.{t}.lora_B"], s * gw @ A.t(), rtol=1e-4, atol=1e-7)
