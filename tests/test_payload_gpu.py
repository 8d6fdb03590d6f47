"""The container entrypoint end to end on one GPU (tiny model): files the controller expects,
dataset ingest from the mount, and resume after a restart landing on the same adapters as an
uninterrupted run."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))


def _args(ckpt, data, steps, extra=()):
    return ["--model=gpt2-tiny", "--seq_len=32", "--batch_size=2", "--epochs=1", f"--steps_per_epoch={steps}",
            "--lr=0.001", "--log_interval=2", "--save_model", f"--dataset_path={data}", f"--checkpoint_path={ckpt}",
            *extra]


def test_entrypoint_outputs_and_resume(tmp_path, capsys):
    from b200ft import train
    data = tmp_path / "dataset"; data.mkdir()
    rng = np.random.default_rng(0)
    rng.integers(0, 503, size=5000, dtype=np.uint16).tofile(data / "tokens.bin")

    a = tmp_path / "a"
    assert train.main(_args(a, data, 6)) == 0
    out = capsys.readouterr().out
    assert out.splitlines()[0].startswith("Epoch"), "the controller's log stream is gated on an 'Epoch' line"
    for f in ("metrics.csv", "adapter.pt", "adapter_config.json", "adapter_model.safetensors", "resume.state"):
        assert (a / f).exists(), f
    import pandas as pd
    rec = pd.read_csv(a / "metrics.csv").fillna(0).to_dict("records")
    assert [r["step"] for r in rec] == [2, 4, 6] and all(r["loss"] > 0 for r in rec)

    # "crash" after 3 steps, restart with the full schedule: resumes at step 4
    b = tmp_path / "b"
    assert train.main(_args(b, data, 3)) == 0
    assert train.main(_args(b, data, 6)) == 0
    assert "resumed at step 3" in capsys.readouterr().out
    wa, wb = torch.load(a / "adapter.pt"), torch.load(b / "adapter.pt")
    for k in wa:   # split-K fp32 atomics: low bits of the adapter gradients depend on arrival order
        d = (wb[k] - wa[k]).abs()
        ok = d <= 1e-6 + 1e-3 * wa[k].abs()
        assert float((~ok).float().mean()) <= 1e-2 and float(d.max()) <= 6 * 0.001, k
    rec_b = pd.read_csv(b / "metrics.csv").fillna(0).to_dict("records")
    assert [r["step"] for r in rec_b][-1] == 6


def test_entrypoint_synthetic_and_clipping(tmp_path):
    from b200ft import train
    c = tmp_path / "c"
    assert train.main(_args(c, tmp_path / "nothing", 2, ["--synthetic", "--max_grad_norm=0.5", "--lora_r=16"])) == 0
    cfg = __import__("json").load(open(c / "adapter_config.json"))
    assert cfg["r"] == 16 and cfg["target_modules"] == ["c_attn"]
    w = torch.load(c / "adapter.pt")
    assert w["h0.c_attn.lora_A"].shape == (64, 16) and w["h1.c_attn.lora_B"].shape == (16, 192)


def test_entrypoint_llama_tiny(tmp_path):
    from b200ft import train
    d = tmp_path / "l"
    args = ["--model=llama-tiny", "--seq_len=32", "--batch_size=2", "--epochs=1", "--steps_per_epoch=3", "--lr=0.001",
            "--log_interval=1", "--save_model", "--synthetic", "--lora_r=16", "--lora_alpha=32",
            f"--dataset_path={tmp_path}", f"--checkpoint_path={d}"]
    assert train.main(args) == 0
    cfg = __import__("json").load(open(d / "adapter_config.json"))
    assert cfg["target_modules"] == ["q_proj", "v_proj"] and cfg["fan_in_fan_out"] is False
    w = torch.load(d / "adapter.pt")
    assert w["l0.q_proj.lora_A"].shape == (16, 64) and w["l1.v_proj.lora_B"].shape == (32, 16)
    assert float(w["l0.v_proj.lora_B"].abs().max()) > 0.0       # PEFT's zero B has moved after 3 steps


def test_entrypoint_full_parameter_mode_and_resume(tmp_path, capsys):
    """--mode=full (BASELINE.json configs[4] on the tiny Llama): model.pt + config.json instead of adapter
    files, loss goes down on a repeated batch schedule, and a restart resumes master weights + moments."""
    from b200ft import train

    def args(ckpt, steps):
        return ["--model=llama-tiny", "--mode=full", "--seq_len=32", "--batch_size=2", "--epochs=1",
                f"--steps_per_epoch={steps}", "--lr=0.003", "--log_interval=1", "--save_model", "--synthetic",
                f"--dataset_path={tmp_path}", f"--checkpoint_path={ckpt}"]

    a, b = tmp_path / "fa", tmp_path / "fb"
    assert train.main(args(a, 6)) == 0
    assert "mode=full" in capsys.readouterr().out.splitlines()[0]
    for f in ("metrics.csv", "model.pt", "config.json", "resume.state"):
        assert (a / f).exists(), f
    assert not (a / "adapter.pt").exists()
    cfg = __import__("json").load(open(a / "config.json"))
    assert cfg["mode"] == "full" and cfg["n_layer"] == 2 and cfg["tie_embeddings"] is True
    w = torch.load(a / "model.pt")
    assert w["l1.down_proj.w"].shape == (64, 128) and w["embed"].shape == (503, 64) and w["norm.g"].dtype == torch.bfloat16
    assert train.main(args(b, 3)) == 0
    assert train.main(args(b, 6)) == 0
    assert "resumed at step 3" in capsys.readouterr().out
    wb = torch.load(b / "model.pt")
    # the embedding / gain gradients are summed with atomics: an element whose gradient is almost zero may move by a
    # different fraction of lr in one of AdamW's early steps.  Nearly all elements agree to bf16 precision.
    for k in w:
        d = (wb[k].float() - w[k].float()).abs()
        assert float((d > 2 ** -7 * 0.1).float().mean()) <= 1e-2 and float(d.max()) <= 6 * 0.003, k


def test_entrypoint_starts_from_pretrained_base_weights(tmp_path):
    """--base_weights: a Hugging Face-named safetensors checkpoint under the dataset mount becomes the frozen
    base.  The first logged loss (PEFT init: B = 0, so the adapters do not act yet) must be the oracle's loss of
    THAT checkpoint on the same batch, not the loss of a random model; the PEFT-named adapter file is written."""
    from safetensors.torch import load_file, save_file
    from b200ft import train
    from oracle import gpt2_lora_ref as ref
    cfg = ref.GPT2_TINY
    p = ref.init_base_params(cfg, seed=11, randomize_all=True)
    p = {k: v * 5.0 if k == "wte" else v for k, v in p.items()}              # loss 6.96: far from a random init's 6.25
    sd = {k: v.detach().clone().contiguous() for k, v in ref.hf_model(cfg, p).state_dict().items() if k != "lm_head.weight"}
    data = tmp_path / "dataset"; (data / "hf").mkdir(parents=True)
    save_file(sd, str(data / "hf" / "model.safetensors"))
    out = tmp_path / "o"
    args = ["--model=gpt2-tiny", "--seq_len=32", "--batch_size=2", "--epochs=1", "--steps_per_epoch=2", "--lr=0.001",
            "--log_interval=1", "--save_model", "--synthetic", "--base_weights=hf", "--save_interval=-1",
            f"--dataset_path={data}", f"--checkpoint_path={out}"]
    assert train.main(args) == 0
    import pandas as pd
    first = pd.read_csv(out / "metrics.csv").to_dict("records")[0]
    ids = torch.randint(0, cfg.vocab, (2, 32), generator=torch.Generator().manual_seed(1234), dtype=torch.int64)
    want = float(ref.causal_lm_loss(ref.forward(p, None, ids, cfg, None), ids))
    rand = float(ref.causal_lm_loss(ref.forward(ref.init_base_params(cfg, 0), None, ids, cfg, None), ids))
    assert abs(first["loss"] - want) / want <= 2e-2 and abs(want - rand) / rand > 5e-2
    assert not (out / "resume.state").exists()                               # --save_interval=-1
    t = load_file(str(out / "adapter_model.safetensors"))
    assert t["base_model.model.transformer.h.0.attn.c_attn.lora_A.weight"].shape == (8, 64)
    w = torch.load(out / "adapter.pt")
    assert torch.equal(t["base_model.model.transformer.h.1.attn.c_attn.lora_B.weight"], w["h1.c_attn.lora_B"].t())
    with pytest.raises(ValueError, match="vocabulary"):                      # ids beyond the model's vocabulary
        bad = tmp_path / "bad"; bad.mkdir()
        np.full(4000, 60000, dtype=np.uint16).tofile(bad / "tokens.bin")
        train.main(["--model=gpt2-tiny", "--seq_len=32", "--batch_size=2", "--steps_per_epoch=1",
                    f"--dataset_path={bad}", f"--checkpoint_path={tmp_path / 'x'}"])
