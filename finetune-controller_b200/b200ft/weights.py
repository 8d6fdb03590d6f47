"""Weight files either side of the step.

IN: pretrained base weights from the dataset mount (the controller's init container copies the uploaded
object there: reference app/jobs/kubeflow/PyTorchJobDeployer.py:70-91) in Hugging Face naming —
``*.safetensors`` (one file, or a directory of shards with or without ``model.safetensors.index.json``)
or a ``torch.save`` state dict — translated to the tensor names the step engines take (the oracle's:
GPT-2 Conv1D weights stay ``[in, out]``, Llama nn.Linear weights stay ``[out, in]``; the engines build
the other orientation themselves with the transpose kernel).

OUT: the trained adapters as ``adapter_model.safetensors`` + ``adapter_config.json`` in the key names and
tensor layouts the PEFT library loads (``base_model.model.<module path>.lora_A.weight`` ``[r, in]``,
``lora_B.weight`` ``[out, r]``; GPT-2's Conv1D target is flagged ``fan_in_fan_out``), so that they match the
plugin's ``store_asset_patterns`` (reference app/models/base/finetuning.py:94-97, extended by
``*.safetensors`` in plugin/b200_lora.py) and are synced by the sidecar (reference PyTorchJobDeployer.py:122-139).
"""
from __future__ import annotations

import glob
import json
import os

import torch

HF_BASE_NAME = {"gpt2-124m": "gpt2", "llama-3.2-1b": "meta-llama/Llama-3.2-1B", "llama-3-8b": "meta-llama/Meta-Llama-3-8B"}


# ------------------------------------------------------------------------------------ reading
def _read_state(path: str) -> dict:
    """{name: tensor} from a .safetensors file, a directory of them, or a torch.save state dict."""
    if os.path.isdir(path):
        files = sorted(glob.glob(os.path.join(path, "*.safetensors")))
        if not files:
            files = sorted(glob.glob(os.path.join(path, "*.pt")) + glob.glob(os.path.join(path, "*.bin")))
        if not files:
            raise FileNotFoundError(f"no *.safetensors / *.pt / *.bin under {path}")
        out = {}
        for f in files:
            if os.path.basename(f).startswith("adapter_"):
                continue
            out.update(_read_state(f))
        return out
    if path.endswith(".safetensors"):
        from safetensors import safe_open
        out = {}
        with safe_open(path, framework="pt", device="cpu") as f:
            for k in f.keys():
                out[k] = f.get_tensor(k)
        return out
    st = torch.load(path, map_location="cpu")
    if isinstance(st, dict) and "state_dict" in st and isinstance(st["state_dict"], dict):
        st = st["state_dict"]
    return {k: v for k, v in st.items() if torch.is_tensor(v)}


def _strip(st: dict, prefixes: tuple) -> dict:
    out = {}
    for k, v in st.items():
        for p in prefixes:
            if k.startswith(p):
                k = k[len(p):]
                break
        out[k] = v
    return out


def gpt2_from_hf(st: dict, dims) -> dict:
    """HF GPT2LMHeadModel names -> engine names.  Conv1D weights are [in, out] on both sides."""
    st = _strip(st, ("transformer.",))
    need = lambda k: st[k].float()  # noqa: E731
    p = {"wte": need("wte.weight"), "wpe": need("wpe.weight"), "lnf.g": need("ln_f.weight"), "lnf.b": need("ln_f.bias")}
    for l in range(dims.n_layer):
        for ours, theirs in (("c_attn", "attn.c_attn"), ("attn_proj", "attn.c_proj"), ("c_fc", "mlp.c_fc"),
                             ("mlp_proj", "mlp.c_proj")):
            p[f"h{l}.{ours}.w"] = need(f"h.{l}.{theirs}.weight")
            p[f"h{l}.{ours}.b"] = need(f"h.{l}.{theirs}.bias")
        for ours, theirs in (("ln1", "ln_1"), ("ln2", "ln_2")):
            p[f"h{l}.{ours}.g"] = need(f"h.{l}.{theirs}.weight")
            p[f"h{l}.{ours}.b"] = need(f"h.{l}.{theirs}.bias")
    return p


def llama_from_hf(st: dict, dims) -> dict:
    """HF LlamaForCausalLM names -> engine names.  nn.Linear weights are [out, in] on both sides."""
    head = st.get("lm_head.weight")
    st = _strip(st, ("model.",))
    need = lambda k: st[k].float()  # noqa: E731
    p = {"embed": need("embed_tokens.weight"), "norm.g": need("norm.weight")}
    for l in range(dims.n_layer):
        for n in ("q_proj", "k_proj", "v_proj", "o_proj"):
            p[f"l{l}.{n}.w"] = need(f"layers.{l}.self_attn.{n}.weight")
        for n in ("gate_proj", "up_proj", "down_proj"):
            p[f"l{l}.{n}.w"] = need(f"layers.{l}.mlp.{n}.weight")
        p[f"l{l}.attn_norm.g"] = need(f"layers.{l}.input_layernorm.weight")
        p[f"l{l}.mlp_norm.g"] = need(f"layers.{l}.post_attention_layernorm.weight")
    if not dims.tie_embeddings:
        if head is None:
            raise KeyError("lm_head.weight missing for a model with untied embeddings")
        p["lm_head.w"] = head.float()
    return p


def _expected_shapes(dims, llama: bool) -> dict:
    d = dims.d_model
    if llama:
        dh, ff = dims.d_head, dims.d_ff
        out = {"embed": (dims.vocab, d), "norm.g": (d,)}
        for l in range(dims.n_layer):
            out.update({f"l{l}.q_proj.w": (dims.n_head * dh, d), f"l{l}.k_proj.w": (dims.n_kv_head * dh, d),
                        f"l{l}.v_proj.w": (dims.n_kv_head * dh, d), f"l{l}.o_proj.w": (d, dims.n_head * dh),
                        f"l{l}.gate_proj.w": (ff, d), f"l{l}.up_proj.w": (ff, d), f"l{l}.down_proj.w": (d, ff),
                        f"l{l}.attn_norm.g": (d,), f"l{l}.mlp_norm.g": (d,)})
        if not dims.tie_embeddings:
            out["lm_head.w"] = (dims.vocab, d)
        return out
    ff = dims.d_ff
    out = {"wte": (dims.vocab, d), "wpe": (dims.ctx, d), "lnf.g": (d,), "lnf.b": (d,)}
    for l in range(dims.n_layer):
        for n, (i, o) in (("c_attn", (d, 3 * d)), ("attn_proj", (d, d)), ("c_fc", (d, ff)), ("mlp_proj", (ff, d))):
            out[f"h{l}.{n}.w"], out[f"h{l}.{n}.b"] = (i, o), (o,)
        for n in ("ln1", "ln2"):
            out[f"h{l}.{n}.g"], out[f"h{l}.{n}.b"] = (d,), (d,)
    return out


def load_base_weights(path: str, dims, llama: bool) -> dict:
    """Base parameters in the engines' naming from `path` (file or directory, HF or engine naming).
    Every tensor's shape is checked against `dims`; a mismatch or a missing tensor raises — a job
    that was asked to start from a checkpoint must not silently train something else."""
    st = _read_state(path)
    want = _expected_shapes(dims, llama)
    if all(k in st for k in want):                       # already in engine naming (model.pt of a full-mode job)
        p = {k: st[k].float() for k in want}
    else:
        try:
            p = llama_from_hf(st, dims) if llama else gpt2_from_hf(st, dims)
        except KeyError as e:
            raise KeyError(f"{path}: tensor {e} not found; this does not look like a "
                           f"{'Llama' if llama else 'GPT-2'} checkpoint") from None
    for k, shape in want.items():
        if tuple(p[k].shape) != tuple(shape):
            raise ValueError(f"{path}: {k} has shape {tuple(p[k].shape)}, the selected model needs {tuple(shape)}")
    return p


# ------------------------------------------------------------------------------------ writing
def peft_adapter_tensors(adapters: dict, llama: bool) -> dict:
    """Engine-named adapters (adapters_state()) -> PEFT key names and layouts.

    PEFT keeps lora_A.weight [r, in] and lora_B.weight [out, r] for every target type; for a Conv1D
    target (GPT-2 c_attn, fan_in_fan_out) the merged delta is (lora_B.weight @ lora_A.weight)^T * scale.
    Engine layouts: GPT-2 lora_A [in, r], lora_B [r, out]; Llama lora_A [r, in], lora_B [out, r]."""
    out = {}
    for k, v in adapters.items():
        layer, mod, which = k.split(".")
        v = v.detach().float().cpu()
        if llama:
            name = f"base_model.model.model.layers.{layer[1:]}.self_attn.{mod}.{which}.weight"
        else:
            name = f"base_model.model.transformer.h.{layer[1:]}.attn.{mod}.{which}.weight"
            v = v.t()
        out[name] = v.contiguous()
    return out


def save_peft_adapter(directory: str, adapters: dict, *, model: str, llama: bool, r: int, alpha: float,
                      steps: int) -> None:
    from safetensors.torch import save_file
    targets = ["q_proj", "v_proj"] if llama else ["c_attn"]
    tmp = os.path.join(directory, "adapter_model.safetensors.tmp")
    save_file(peft_adapter_tensors(adapters, llama), tmp, metadata={"format": "pt"})
    os.replace(tmp, os.path.join(directory, "adapter_model.safetensors"))
    cfg = {"peft_type": "LORA", "task_type": "CAUSAL_LM", "base_model_name_or_path": HF_BASE_NAME.get(model, model),
           "r": r, "lora_alpha": alpha, "lora_dropout": 0.0, "target_modules": targets, "bias": "none",
           "fan_in_fan_out": not llama, "inference_mode": True, "init_lora_weights": True,
           "modules_to_save": None, "use_rslora": False, "use_dora": False,
           # not PEFT fields (ignored by its loader): provenance of the job
           "base_model": model, "steps": steps}
    with open(os.path.join(directory, "adapter_config.json"), "w") as f:
        json.dump(cfg, f, indent=1)
