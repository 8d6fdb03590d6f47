"""BASELINE.json configs[2] and [3] at their FULL sizes (Llama-3.2-1B, 4 x 2048 tokens; Llama-3-8B, 1 x 4096 tokens),
where the CPU oracle would need minutes per step, through size-independent properties:

  * the loss of a randomly initialised model is ln(vocab) plus half the logit variance of the N(0, 0.02) init
    (0.02^2 * d_model / 2 = 0.4 at d = 2048, 0.8 at d = 4096): within 10 % of ln(vocab);
  * INDEPENDENT attention implementations — this library's tcgen05 kernels, cuDNN's and the flash-attention kernels
    torch SDPA carries — give the same loss and the same adapter gradients for the same weights and batch; everything
    else in the steps is identical code, so this isolates attention forward + backward at full size.  A fixed
    tolerance does not exist for that comparison: a 16-32 layer bf16 model at random init amplifies one-ulp
    differences of the attention output (measured on B200: cuDNN against flash 4 % relative L2 at Llama-3.2-1B,
    13 % at Llama-3-8B; own against either 4 % / 11-12 %).  So the yardstick is measured in the test: the worst
    relative L2 between the two LIBRARY implementations; own kernels against cuDNN must stay within 1.25 x that,
    with cosine >= 0.99 for every tensor;
  * a second step on the same batch lowers the loss (the update has the right sign), and stays finite.

The exact-parity tests against the oracle at these models' widths run on one layer (tests/test_dims8b_gpu.py,
tests/test_llama_gpu.py)."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _run(model, batch, seq, r, attention, backend=None):
    from b200ft import gpt2, params
    if backend is not None:
        from torch.nn.attention import sdpa_kernel
        with sdpa_kernel(backend):
            return _run(model, batch, seq, r, attention)
    eng, dims = params.build_engine(model, gpt2.LoRASpec(r, 2.0 * r), gpt2.OptSpec(lr=1e-3), batch=batch, seq=seq,
                                    device=torch.device(DEV), seed=0, randomize_b=True, attention=attention)
    assert eng.attention == attention
    ids = torch.randint(0, dims.vocab, (batch, seq), generator=torch.Generator().manual_seed(5), dtype=torch.int64).to(DEV)
    l1 = float(eng.step(ids, ids).cpu())
    grads = {k: v.clone() for k, v in eng.grads_state().items()}
    l2 = float(eng.step(ids, ids).cpu())
    vocab = dims.vocab
    del eng
    torch.cuda.empty_cache()
    return l1, l2, grads, vocab


def _worst(ga, gb):
    """(worst relative L2, worst cosine, tensor name of the worst relative L2) over the adapter gradients."""
    rel_w, cos_w, name = 0.0, 1.0, ""
    for k, a in ga.items():
        a64, b64 = a.flatten().double(), gb[k].flatten().double()
        cos = float((a64 @ b64) / (a64.norm() * b64.norm() + 1e-30))
        rel = float((a64 - b64).norm() / (b64.norm() + 1e-30))
        cos_w = min(cos_w, cos)
        if rel > rel_w:
            rel_w, name = rel, k
    return rel_w, cos_w, name


@pytest.mark.parametrize("model,batch,seq,r", [("llama-3.2-1b", 4, 2048, 16), ("llama-3-8b", 1, 4096, 16)])
def test_full_size_llama_lora_step_own_attention_agrees_with_library_kernels(model, batch, seq, r):
    from torch.nn.attention import SDPBackend
    own = _run(model, batch, seq, r, "own")
    lib = _run(model, batch, seq, r, "sdpa", SDPBackend.CUDNN_ATTENTION)
    alt = _run(model, batch, seq, r, "sdpa", SDPBackend.FLASH_ATTENTION)
    l1, l2, g_own, vocab = own
    assert 0.0 <= l1 - math.log(vocab) <= 0.10 * math.log(vocab), (l1, math.log(vocab))
    assert math.isfinite(l2) and l2 < l1, (l1, l2)
    for other in (lib, alt):
        assert abs(l1 - other[0]) <= 1e-3 * l1 and abs(l2 - other[1]) <= 1e-3 * l1, (own[:2], other[:2])
    yard, _, _ = _worst(lib[2], alt[2])
    rel, cos, name = _worst(g_own, lib[2])
    assert all(torch.isfinite(g).all() for g in g_own.values())
    assert 0.0 < yard <= 0.25, f"library kernels disagree by {yard}: the yardstick itself is off"
    assert rel <= 1.25 * yard and cos >= 0.99, f"{name}: rel {rel} (cuDNN vs flash: {yard}), worst cos {cos}"
