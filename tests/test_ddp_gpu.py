"""Two ranks over NCCL on two GPUs: the bucketed gradient all-reduce inside backward plus the
1/world scale folded into AdamW must reproduce the oracle's data-parallel emulation (mean of
the per-rank gradients), and both ranks must end with identical adapters.
Skipped on a single-GPU box."""
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "finetune-controller_b200")


def _rank_llama(rank, world, port, out_dir):
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist
    from b200ft import ddp, gpt2, llama
    from oracle import llama_lora_ref as ref
    from oracle.gpt2_lora_ref import AdamWConfig
    ddp.init_process_group_from_env("nccl")
    cfg, lora, opt = ref.LLAMA_TINY, ref.LoRAConfig(r=16, alpha=32.0), AdamWConfig(lr=1e-3)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = ref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    dims = llama.LlamaDims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.n_kv_head, cfg.d_ff, cfg.vocab, cfg.ctx,
                           cfg.rms_eps, cfg.rope_theta, cfg.rope_llama3, tie_embeddings=cfg.tie_embeddings)
    eng = llama.LlamaLoRAStep(dims, gpt2.LoRASpec(lora.r, lora.alpha),
                              gpt2.OptSpec(opt.lr, opt.beta1, opt.beta2, opt.eps, opt.weight_decay), base, ad,
                              batch=2, seq=32, device=f"cuda:{rank}", world_size=world,
                              grad_sync=ddp.BucketedGradSync(cfg.n_layer, n_buckets=2))
    ids, labels = ref.synthetic_batch(cfg, 2, 32, seed=77, rank=rank)
    loss = float(eng.step(ids.cuda(rank), labels.cuda(rank)).cpu())
    torch.save({"loss": loss, "adapters": eng.adapters_state()}, os.path.join(out_dir, f"l{rank}.pt"))
    dist.destroy_process_group()


def test_llama_two_ranks_stay_identical(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from b200ft import ddp
    mp.spawn(_rank_llama, args=(2, ddp.free_port(), str(tmp_path)), nprocs=2, join=True)
    a, b = torch.load(tmp_path / "l0.pt"), torch.load(tmp_path / "l1.pt")
    assert a["loss"] != b["loss"]                      # different micro-batches
    for k in a["adapters"]:
        assert torch.equal(a["adapters"][k], b["adapters"][k]), k


def _rank(rank, world, port, out_dir):
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist
    from b200ft import ddp, gpt2
    from oracle import gpt2_lora_ref as ref
    ddp.init_process_group_from_env("nccl")
    cfg, lora, opt = ref.GPT2_TINY, ref.LoRAConfig(r=8, alpha=16.0), ref.AdamWConfig(lr=1e-3)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = ref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    dims = gpt2.GPT2Dims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.d_ff, cfg.vocab, cfg.ctx, cfg.ln_eps)
    sync = ddp.BucketedGradSync(cfg.n_layer, n_buckets=2)
    eng = gpt2.GPT2LoRAStep(dims, gpt2.LoRASpec(lora.r, lora.alpha),
                            gpt2.OptSpec(opt.lr, opt.beta1, opt.beta2, opt.eps, opt.weight_decay),
                            base, ad, batch=2, seq=32, device=f"cuda:{rank}", world_size=world, grad_sync=sync)
    losses = []
    for s in range(2):
        ids, labels = ref.synthetic_batch(cfg, 2, 32, seed=1234 + 10 * s, rank=rank)
        losses.append(float(eng.step(ids.cuda(rank), labels.cuda(rank)).cpu()))
    torch.save({"losses": losses, "adapters": eng.adapters_state(), "grads": eng.grads_state()},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def test_two_rank_step_matches_oracle_data_parallel(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from b200ft import ddp
    from oracle import gpt2_lora_ref as ref
    world = 2
    mp.spawn(_rank, args=(world, ddp.free_port(), str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    cfg, lora, opt = ref.GPT2_TINY, ref.LoRAConfig(r=8, alpha=16.0), ref.AdamWConfig(lr=1e-3)
    ad = ref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    tr = ref.OracleTrainer(cfg, lora, opt, ref.init_base_params(cfg, seed=0, randomize_all=True), ad)
    for s in range(2):
        batches = [ref.synthetic_batch(cfg, 2, 32, seed=1234 + 10 * s, rank=r) for r in range(world)]
        mean_loss = tr.step(None, None, world_grads=batches)
        got = sum(o["losses"][s] for o in outs) / world
        assert abs(got - mean_loss) / mean_loss <= 2e-2
    for k, t in tr.trainable.items():
        a0, a1 = outs[0]["adapters"][k], outs[1]["adapters"][k]
        assert torch.equal(a0, a1), f"ranks diverged on {k}"
        d_ref, d_gpu = (t.detach() - ad[k]).flatten().double(), (a0 - ad[k]).flatten().double()
        cos = float(d_gpu @ d_ref / (d_gpu.norm() * d_ref.norm() + 1e-30))
        assert cos >= 0.99, (k, cos)
        # the reduced gradient buffers hold the SUM over ranks (1/world lives in AdamW's grad_scale)
        g = outs[0]["grads"][k].flatten().double() / world
        gr = tr.last_grads[k].flatten().double()
        assert float(g @ gr / (g.norm() * gr.norm() + 1e-30)) >= 0.99


def _rank_full(rank, world, port, out_dir):
    for p in (ROOT, PKG):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    import torch.distributed as dist
    from b200ft import ddp, gpt2, llama, llama_full
    from oracle import llama_lora_ref as ref
    from oracle.gpt2_lora_ref import AdamWConfig
    ddp.init_process_group_from_env("nccl")
    cfg, opt = _full_cfg(ref), AdamWConfig(lr=1e-3, max_grad_norm=1.0)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    dims = llama.LlamaDims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.n_kv_head, cfg.d_ff, cfg.vocab, cfg.ctx,
                           cfg.rms_eps, cfg.rope_theta, cfg.rope_llama3, tie_embeddings=cfg.tie_embeddings)
    eng = llama_full.LlamaFullStep(dims, gpt2.OptSpec(opt.lr, opt.beta1, opt.beta2, opt.eps, opt.weight_decay,
                                                      opt.max_grad_norm), base, batch=2, seq=32,
                                   device=f"cuda:{rank}", world_size=world,
                                   # the production configuration of bench.py / train.py: per-layer buckets on a
                                   # CTA-limited communicator, SMs reserved in the backward GEMMs, tail on the default group
                                   grad_sync=ddp.BucketedGradSync(cfg.n_layer, n_buckets=cfg.n_layer,
                                                                  group=ddp.overlap_group(8), reserve_ctas=8,
                                                                  use_tail_group=True))
    assert eng.bwd_ctas == 148 - 16
    losses = []
    for s in range(2):
        ids, labels = ref.synthetic_batch(cfg, 2, 32, seed=500 + 10 * s, rank=rank)
        losses.append(float(eng.step(ids.cuda(rank), labels.cuda(rank)).cpu()))
    torch.save({"losses": losses, "params": eng.params_state()}, os.path.join(out_dir, f"f{rank}.pt"))
    dist.destroy_process_group()


def _full_cfg(ref):
    return ref.LlamaConfig(n_layer=3, d_model=64, n_head=4, n_kv_head=2, d_ff=128, vocab=503, ctx=128,
                           rope_theta=10000.0, rope_llama3=False, tie_embeddings=False)


def test_full_parameter_two_ranks_match_oracle_data_parallel(tmp_path):
    """Full-parameter step, one all-reduce bucket per layer (+ lm_head on the top bucket, embedding table
    and norm gains on the bottom one), global-norm clipping on the reduced gradients."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    from b200ft import ddp
    from oracle import llama_lora_ref as ref
    from oracle.gpt2_lora_ref import AdamWConfig
    world = 2
    mp.spawn(_rank_full, args=(world, ddp.free_port(), str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(tmp_path / f"f{r}.pt") for r in range(world)]
    cfg = _full_cfg(ref)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    tr = ref.FullOracleTrainer(cfg, AdamWConfig(lr=1e-3, max_grad_norm=1.0), base)
    for s in range(2):
        batches = [ref.synthetic_batch(cfg, 2, 32, seed=500 + 10 * s, rank=r) for r in range(world)]
        mean_loss = tr.step(None, None, world_grads=batches)
        got = sum(o["losses"][s] for o in outs) / world
        assert abs(got - mean_loss) / mean_loss <= 2e-2
    for k, t in tr.trainable.items():
        a0, a1 = outs[0]["params"][k], outs[1]["params"][k]
        assert torch.equal(a0, a1), f"ranks diverged on {k}"
        d_ref, d_gpu = (t.detach() - base[k]).flatten().double(), (a0 - base[k]).flatten().double()
        cos = float(d_gpu @ d_ref / (d_gpu.norm() * d_ref.norm() + 1e-30))
        assert cos >= 0.97, (k, cos)
