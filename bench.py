#!/usr/bin/env python
"""Benchmark of the fine-tuning step (BASELINE.json metric: fine-tune tokens/sec + % roofline,
next to the CPU reference step).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N=1 workload = BASELINE.json configs[1]: GPT-2 124M LoRA r=8 bf16, seq 1024, batch 32 per GPU.
N>1 (launched by torchrun, one rank per GPU): the same per-GPU batch on every rank (weak
scaling), gradients all-reduced over NCCL.  Rank 0 prints ONE JSON line.

  value     tokens/s, whole job, token ids already resident in HBM, CUDA-event timed, max over ranks
  e2e       tokens/s through the public call GPT2LoRAStep.step_host(): per step the ids/labels
            are copied from pinned HOST memory and the loss is read back to the host
  roofline  the dominant kernel (tcgen05 GEMM): algorithmic FLOPs of every GEMM launch of a step
            / the CUDA-event duration of those launches, measured in an instrumented pass of
            the same step right after the timed region (the events would perturb `value`)
  cpu_baseline  the oracle's step (plain torch fp32, CPU) at BASELINE.json configs[0], timed
            on this box's host cores in the same run (N=1, rank 0 only)

--impl reference times the CPU restatement of the reference step (oracle/) on the host cores,
each "step" a bounded sample (1 sequence of seq 1024) of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "finetune-controller_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

# DRAM bytes (read + write) per launch of the dominant kernel, averaged over the 98 launches of
# one step: ncu `dram__bytes_read.sum + dram__bytes_write.sum` (see GEMM_DRAM_SOURCE).  None until
# the capture for the current kernel has been taken.
GEMM_DRAM_BYTES_PER_LAUNCH = 351.8e6
GEMM_DRAM_SOURCE = ("profiles/r02_step_launch_summary_c2.txt: 34.47 GB over the 98 launches of one step "
                    "(ncu dram__bytes_read.sum + dram__bytes_write.sum); algorithmic operand + output bytes of "
                    "those launches: 32.5 GB")

METRIC = "fine-tune tokens/sec"
UNIT = "tokens/s"
MODEL = "gpt2-124m"
BATCH, SEQ, LORA_R, LORA_ALPHA = 32, 1024, 8, 16.0
# name -> (model, batch per GPU, seq, lora r, lora alpha, description).  The driver's runs use the
# default (c2 = BASELINE.json configs[1]); c3 / c4 are the Llama configs, for our own measurements.
WORKLOADS = {
    "c2": ("gpt2-124m", 32, 1024, 8, 16.0, "GPT-2 124M LoRA r=8 bf16 fine-tune step (fwd+loss+bwd+AdamW), seq 1024, "
                                          "batch 32/GPU (BASELINE.json configs[1])"),
    "c3": ("llama-3.2-1b", 4, 2048, 16, 32.0, "Llama-3.2-1B LoRA r=16 (q_proj, v_proj) bf16 fine-tune step, seq 2048, "
                                              "batch 4/GPU (BASELINE.json configs[2])"),
    "c4": ("llama-3-8b", 1, 4096, 16, 32.0, "Llama-3-8B LoRA r=16 (q_proj, v_proj) bf16 fine-tune step, seq 4096, "
                                            "batch 1/GPU (BASELINE.json configs[3])"),
    "c5": ("llama-3-8b", 1, 4096, 0, 0.0, "Llama-3-8B FULL-parameter bf16 fine-tune step (fp32 master + Adam moments, "
                                          "fused AdamW over 8.03 B parameters), seq 4096, batch 1/GPU "
                                          "(BASELINE.json configs[4])"),
    "c5s": ("llama-3.2-1b", 2, 2048, 0, 0.0, "Llama-3.2-1B FULL-parameter bf16 fine-tune step, seq 2048, batch 2/GPU "
                                             "(small sibling of BASELINE.json configs[4])"),
}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        return {"hbm_gbs": d["hbm_gbs"], "tf_burst": d["bf16_tflops"], "tf_sustained": d["bf16_tflops_sustained"],
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "tf_burst": 1590.0, "tf_sustained": 1400.0, "source": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.rows, self.proc = gpu_index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[4:8]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------
# CPU arms (the oracle is only ever the thing compared against / reported beside)
# ----------------------------------------------------------------------------------------
def cpu_oracle_tokens_per_s(batch: int, seq: int, steps: int, warmup: int) -> tuple[float, int, float]:
    from oracle import gpt2_lora_ref as ref
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg, lora, opt = ref.GPT2_124M, ref.LoRAConfig(r=LORA_R, alpha=LORA_ALPHA), ref.AdamWConfig()
    base = ref.init_base_params(cfg, seed=0)
    ad = ref.init_lora_params(cfg, lora, seed=1)
    tr = ref.OracleTrainer(cfg, lora, opt, base, ad)
    times = []
    for s in range(warmup + steps):
        ids, labels = ref.synthetic_batch(cfg, batch, seq, seed=1234 + s)
        t0 = time.perf_counter()
        tr.step(ids, labels)
        dt = time.perf_counter() - t0
        if s >= warmup:
            times.append(dt)
    per_step = statistics.median(times)
    return batch * seq / per_step, torch.get_num_threads(), per_step


def run_reference(args) -> None:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    sample_b, sample_t = 1, SEQ
    tps, cores, per_step = cpu_oracle_tokens_per_s(sample_b, sample_t, args.steps, args.warmup)
    sample = (f"oracle/gpt2_lora_ref.py (CPU restatement of the reference step; the reference ships no "
              f"compute payload), GPT-2 124M LoRA r={LORA_R} fp32, each step = {sample_b} sequence of {sample_t} "
              f"tokens of the C2 workload, {cores} host threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": tps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"GPT-2 124M LoRA r={LORA_R} fine-tune step, seq {SEQ}, batch {BATCH}/GPU (C2)",
                   "sample_per_step": f"{sample_b}x{sample_t} tokens"},
        "cpu_baseline": {"value": tps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": tps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------
# multi-rank parity check (N > 1 only; runs BEFORE the timed region): the data-parallel step on the real NCCL
# ranks against the oracle's data-parallel emulation (mean of the per-rank gradients), for the two exchange
# paths the payload uses — LoRA (two buckets on the default communicator) and full-parameter (one bucket per
# layer on the CTA-limited communicator, SMs reserved in the backward GEMMs, tail on the default group).
# The GPU test box has one GPU, so tests/test_ddp_gpu.py can only run where 2+ GPUs are leased; this puts the
# same check in front of every multi-GPU bench line.
# ----------------------------------------------------------------------------------------
def dp_parity(rank: int, world: int, dev, nccl_ctas: int) -> dict:
    import torch.distributed as dist
    from b200ft import ddp, gpt2, llama, llama_full
    from oracle import gpt2_lora_ref as gref
    from oracle import llama_lora_ref as lref

    def cos(a, b):
        a, b = a.flatten().double(), b.flatten().double()
        return float((a @ b) / (a.norm() * b.norm() + 1e-30))

    def identical(flat: torch.Tensor) -> bool:
        """bit-equality of a rank's trained parameters with rank 0's"""
        mine = flat.detach().clone()
        dist.broadcast(flat0 := mine.clone(), src=0)
        same = torch.tensor([float(torch.equal(mine, flat0))], device=dev)
        dist.all_reduce(same, op=dist.ReduceOp.MIN)
        return bool(same.item() == 1.0)

    def mean_loss(l: torch.Tensor) -> float:
        t = l.detach().clone().float()
        dist.all_reduce(t)
        return float(t.item()) / world

    out, steps = {}, 2
    # --- LoRA path (GPT-2 tiny) ---
    cfg, lora, opt = gref.GPT2_TINY, gref.LoRAConfig(r=8, alpha=16.0), gref.AdamWConfig(lr=1e-3)
    base = gref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = gref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    dims = gpt2.GPT2Dims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.d_ff, cfg.vocab, cfg.ctx, cfg.ln_eps)
    eng = gpt2.GPT2LoRAStep(dims, gpt2.LoRASpec(8, 16.0), gpt2.OptSpec(opt.lr, opt.beta1, opt.beta2, opt.eps,
                                                                     opt.weight_decay), base, ad, batch=2, seq=32,
                            device=dev, world_size=world, grad_sync=ddp.BucketedGradSync(cfg.n_layer, n_buckets=2))
    got = []
    for s in range(steps):
        ids, labels = gref.synthetic_batch(cfg, 2, 32, seed=1234 + 10 * s, rank=rank)
        got.append(mean_loss(eng.step(ids.to(dev), labels.to(dev))))
    same_lora = identical(eng.P)
    rel_lora, cos_lora = 0.0, 1.0
    if rank == 0:
        tr = gref.OracleTrainer(cfg, lora, opt, base, ad)
        for s in range(steps):
            want = tr.step(None, None, world_grads=[gref.synthetic_batch(cfg, 2, 32, seed=1234 + 10 * s, rank=r)
                                                    for r in range(world)])
            rel_lora = max(rel_lora, abs(got[s] - want) / want)
        new = eng.adapters_state()
        cos_lora = min(cos(new[k] - ad[k], t.detach() - ad[k]) for k, t in tr.trainable.items())
    # --- full-parameter path (Llama tiny, 3 layers, untied), production all-reduce configuration ---
    cfg = lref.LlamaConfig(n_layer=3, d_model=64, n_head=4, n_kv_head=2, d_ff=128, vocab=503, ctx=128,
                           rope_theta=10000.0, rope_llama3=False, tie_embeddings=False)
    opt = gref.AdamWConfig(lr=1e-3, max_grad_norm=1.0)
    base = lref.init_base_params(cfg, seed=0, randomize_all=True)
    ldims = llama.LlamaDims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.n_kv_head, cfg.d_ff, cfg.vocab, cfg.ctx,
                            cfg.rms_eps, cfg.rope_theta, cfg.rope_llama3, tie_embeddings=cfg.tie_embeddings)
    k = nccl_ctas if nccl_ctas > 0 else 8
    sync = ddp.BucketedGradSync(cfg.n_layer, n_buckets=cfg.n_layer, group=ddp.overlap_group(k), reserve_ctas=k,
                                use_tail_group=True)
    feng = llama_full.LlamaFullStep(ldims, gpt2.OptSpec(opt.lr, opt.beta1, opt.beta2, opt.eps, opt.weight_decay,
                                                        opt.max_grad_norm), base, batch=2, seq=32, device=dev,
                                    world_size=world, grad_sync=sync)
    got = []
    for s in range(steps):
        ids, labels = lref.synthetic_batch(cfg, 2, 32, seed=500 + 10 * s, rank=rank)
        got.append(mean_loss(feng.step(ids.to(dev), labels.to(dev))))
    same_full = identical(feng.P) and identical(feng.Pn)
    rel_full, cos_full = 0.0, 1.0
    if rank == 0:
        tr = lref.FullOracleTrainer(cfg, opt, base)
        for s in range(steps):
            want = tr.step(None, None, world_grads=[lref.synthetic_batch(cfg, 2, 32, seed=500 + 10 * s, rank=r)
                                                    for r in range(world)])
            rel_full = max(rel_full, abs(got[s] - want) / want)
        new = feng.params_state()
        cos_full = min(cos(new[k2] - base[k2], t.detach() - base[k2]) for k2, t in tr.trainable.items())
    del eng, feng
    torch.cuda.empty_cache()
    out = {"ok": bool(same_lora and same_full and rel_lora <= 2e-2 and rel_full <= 2e-2 and cos_lora >= 0.99
                      and cos_full >= 0.97),
           "loss_rel": max(rel_lora, rel_full), "ranks_identical": bool(same_lora and same_full),
           "lora": {"loss_rel": rel_lora, "delta_cos_min": cos_lora, "ranks_identical": same_lora},
           "full": {"loss_rel": rel_full, "delta_cos_min": cos_full, "ranks_identical": same_full,
                    "nccl_ctas": k},
           "what": f"{steps} steps of tiny GPT-2 LoRA and tiny Llama full-parameter on the {world} NCCL ranks vs the CPU "
                   "oracle's data-parallel emulation; tolerances: loss 2e-2 rel, parameter-delta cosine 0.99 / 0.97, "
                   "bit-identical parameters across ranks"}
    return out


# ----------------------------------------------------------------------------------------
# the CUDA arm
# ----------------------------------------------------------------------------------------
def run_ours(args) -> None:
    import torch.distributed as dist
    from b200ft import ddp, gpt2, ops, params

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 path has no CPU fallback "
                         "(use --impl reference for the CPU arm)")
    rank, world, local = ddp.init_process_group_from_env()
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    model, BATCH, SEQ, lora_r, lora_alpha, workload_desc = WORKLOADS[args.workload]
    full = lora_r == 0
    parity = dp_parity(rank, world, dev, args.nccl_ctas) if world > 1 and not args.no_dp_parity else None
    lora = gpt2.LoRASpec(lora_r or 8, lora_alpha or 16.0)
    opt = gpt2.OptSpec()
    n_layer = params.MODELS[model].n_layer
    # LoRA gradients are a few MB: two buckets.  Full-parameter gradients are hundreds of MB per layer:
    # one bucket per layer, each reduced while the layers below it are still in backward
    sync = None
    if world > 1 and full:
        k = args.nccl_ctas
        sync = ddp.BucketedGradSync(n_layer, n_buckets=n_layer, group=ddp.overlap_group(k) if k > 0 else None,
                                    reserve_ctas=k, use_tail_group=True, tail_layers=args.tail_layers)
    elif world > 1:
        sync = ddp.BucketedGradSync(n_layer, n_buckets=1)   # the whole flat adapter gradient: ONE all-reduce per step
    # random-init weights; B drawn at random (not PEFT's zero) so every gradient path carries data
    eng, dims = params.build_engine(model, lora, opt, batch=BATCH, seq=SEQ, device=dev, seed=0, world_size=world,
                                    grad_sync=sync, randomize_b=True, mode="full" if full else "lora",
                                    attention=args.attention, overlap_optimizer=args.overlap_optimizer)
    g = torch.Generator().manual_seed(1234 + rank)
    n_sets = 4
    host = [torch.randint(0, dims.vocab, (BATCH, SEQ), generator=g, dtype=torch.int64).pin_memory()
            for _ in range(n_sets)]
    resident = [h.to(dev) for h in host]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for s in range(steps):
            fn(s)
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def step_resident(s):
        eng.step(resident[s % n_sets], resident[s % n_sets])

    losses = []

    def step_host(s):
        losses.append(eng.step_host(host[s % n_sets], host[s % n_sets]))

    for s in range(args.warmup):
        step_resident(s)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = ops.LAUNCHES
    bytes0 = sync.bytes_reduced if sync is not None else 0
    ms = timed(step_resident, args.steps)
    launches = ops.LAUNCHES - launches0
    exchanged = (sync.bytes_reduced - bytes0) // args.steps if sync is not None else 0
    ms_e2e = timed(step_host, args.steps)
    clocks = sampler.stop() if rank == 0 else None
    tokens = world * BATCH * SEQ * args.steps

    # instrumented pass: per-kernel-class CUDA events inside the same step
    prof_steps = min(args.steps, 5)
    ops.PROFILE = {}
    for s in range(prof_steps):
        step_resident(s)
    torch.cuda.synchronize()
    prof, ops.PROFILE = ops.PROFILE, None
    pk = peaks()
    # N > 1: what the gradient exchange costs the step — the same timed loop with the exchange switched off (the ranks
    # then drift apart numerically, so this runs last, after every number that is reported as a result)
    exchange = None
    if world > 1 and sync is not None:
        eng.grad_sync = None
        step_resident(0)
        ms_local = timed(step_resident, args.steps)
        eng.grad_sync = sync
        exchange = {"ms_per_step_without_exchange": ms_local / args.steps,
                    "exposed_exchange_ms": (ms - ms_local) / args.steps,
                    "all_reduces_per_step": len(sync.buckets) if not full else None,
                    "gradient_bytes_per_step": exchanged,
                    "note": "difference of two timed loops of the same steps (max over ranks each); includes the skew "
                            "between ranks that the all-reduce makes visible"}
    classes = {}
    for name, evs in prof.items():
        t_ms = sum(e[0].elapsed_time(e[1]) for e in evs)
        work = sum(e[2] for e in evs)
        classes[name] = {"launches_per_step": len(evs) / prof_steps, "ms_per_step": t_ms / prof_steps,
                         "work_per_step": work / prof_steps}
    shapes = {}
    for e in prof["gemm"]:
        rec = shapes.setdefault(e[3], [0, 0.0, 0.0])
        rec[0] += 1; rec[1] += e[0].elapsed_time(e[1]); rec[2] += e[2]
    gemm_shapes = [{"m_n_k_amn_bmn_epi_cg": list(k), "launches_per_step": v[0] / prof_steps,
                    "us_per_launch": 1e3 * v[1] / v[0], "tflops": v[2] / (v[1] * 1e-3) / 1e12}
                   for k, v in sorted(shapes.items(), key=lambda kv: -kv[1][1])]
    # dominant kernel = the cta_group::2 instantiation (every large GEMM of the step); the N = 16
    # adapter GEMMs and the split-K adapter-gradient contractions run the <1> instantiation
    big = [e for e in prof["gemm"] if e[3][6] == 2]
    gm = {"launches_per_step": len(big) / prof_steps,
          "ms_per_step": sum(e[0].elapsed_time(e[1]) for e in big) / prof_steps,
          "work_per_step": sum(e[2] for e in big) / prof_steps}
    tf = gm["work_per_step"] / (gm["ms_per_step"] * 1e-3) / 1e12
    roofline = {"kernel": "gemm_bf16_tcgen05_kernel<2>", "bound": "tensor", "achieved": tf, "peak": pk["tf_sustained"],
                "unit": "TFLOP/s", "frac": tf / pk["tf_sustained"],
                "flop_per_launch": gm["work_per_step"] / gm["launches_per_step"],
                "us_per_launch": 1e3 * gm["ms_per_step"] / gm["launches_per_step"],
                "traffic": GEMM_DRAM_BYTES_PER_LAUNCH if args.workload == "c2" else None,
                "traffic_source": GEMM_DRAM_SOURCE if args.workload == "c2" else None,
                "peak_source": pk["source"] + ", sustained bf16 (kernel timed inside a long step)",
                "launches_per_step": gm["launches_per_step"], "ms_per_step": gm["ms_per_step"],
                "share_of_step": gm["ms_per_step"] / (ms / args.steps)}
    others = []
    for name in ("softmax_xent", "adamw", "layernorm_fwd", "layernorm_bwd", "rmsnorm_fwd", "rmsnorm_bwd", "rope",
                 "swiglu_fwd", "swiglu_bwd", "lora_grad", "rmsnorm_dgamma", "embed_bwd"):
        if name in classes:
            c = classes[name]
            gbs = c["work_per_step"] / (c["ms_per_step"] * 1e-3) / 1e9
            others.append({"kernel": name, "bound": "hbm", "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s",
                           "frac": gbs / pk["hbm_gbs"], "ms_per_step": c["ms_per_step"],
                           "launches_per_step": c["launches_per_step"]})

    # N > 1: the configurations whose exchange is not negligible, measured briefly after the headline's timed region
    # (C3: Llama-3.2-1B LoRA, 3.4 MB of gradients per step; C5: Llama-3-8B full-parameter, 16 GB per step)
    extras = None
    attention_used = eng.attention
    if world > 1 and args.workload == "c2" and not args.no_extra_configs:
        del eng, resident
        torch.cuda.empty_cache()
        extras = {}
        for name in ("c3", "c5"):
            extras[name] = extra_config(name, args, rank, world, dev, pk)
            torch.cuda.empty_cache()
    if rank != 0:
        return
    if args.detail:
        with open(args.detail, "w") as f:
            json.dump({"gemm_shapes": gemm_shapes, "classes": classes}, f, indent=1)
    # north_star asks for the fused AdamW's HBM fraction; the LoRA parameter set (1.2 M) is
    # launch-latency-bound, so it is also timed on a 256 Mi-element buffer
    if world == 1 and not full:
        others.append(adamw_large_buffer(ops, dev, pk))
    cpu = None
    if world == 1 and not args.no_cpu_baseline and args.workload == "c2":
        tps, cores, per_step = cpu_oracle_tokens_per_s(8, 512, steps=3, warmup=1)
        cpu = {"value": tps, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": "oracle/gpt2_lora_ref.py at BASELINE.json configs[0] (GPT-2 124M LoRA r=8 fp32, batch 8, "
                         f"seq 512): median of 3 steps after 1 warm-up (BASELINE.md §4), {per_step:.2f} s/step"}
    line = {
        "metric": METRIC, "value": tokens / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload_desc,
                   "global_batch": BATCH * world, "seq_len": SEQ, "parallelism": f"dp{world}",
                   "l2": "per-step working set (GBs of activations and logits) is far larger than the "
                         "126 MB L2; no explicit flush",
                   "attention": ("torch SDPA (library); every other op is this repo's CUDA" if attention_used == "sdpa"
                                 else "this repo's tcgen05 flash-attention forward + backward (no library kernel in the step)")},
        "e2e": {"value": tokens / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                "h2d_bytes_per_step": 2 * BATCH * SEQ * 8, "d2h_bytes_per_step": 4, "last_loss": losses[-1]},
        "gpu_launches": launches,
        "clocks": clocks,
        "roofline": roofline,
        "roofline_others": others,
        "cpu_baseline": cpu,
    }
    if parity is not None:
        line["dp_parity"] = parity
    if exchange is not None:
        line["exchange"] = exchange
    if extras is not None:
        line["configs_extra"] = extras
    print(json.dumps(line), flush=True)


def extra_config(name: str, args, rank: int, world: int, dev, pk) -> dict:
    """A short run (3 warm-up + 5 timed steps, twice: with the gradient exchange and with it switched off) of another
    BASELINE.json configuration on the same ranks: ms/step, tokens/s, GEMM TFLOP/s inside the step and the part of the
    exchange that backward does not hide (difference of the two timings)."""
    import torch.distributed as dist
    from b200ft import ddp, gpt2, ops, params
    model, batch, seq, lora_r, lora_alpha, desc = WORKLOADS[name]
    full = lora_r == 0
    n_layer = params.MODELS[model].n_layer
    try:
        if full:
            k = args.nccl_ctas
            sync = ddp.BucketedGradSync(n_layer, n_buckets=n_layer, group=ddp.overlap_group(k) if k > 0 else None,
                                        reserve_ctas=k, use_tail_group=True)
        else:
            sync = ddp.BucketedGradSync(n_layer, n_buckets=1)
        eng, dims = params.build_engine(model, gpt2.LoRASpec(lora_r or 8, lora_alpha or 16.0), gpt2.OptSpec(), batch=batch,
                                        seq=seq, device=dev, seed=0, world_size=world, grad_sync=sync, randomize_b=True,
                                        mode="full" if full else "lora", attention=args.attention)
        g = torch.Generator().manual_seed(4321 + rank)
        ids = [torch.randint(0, dims.vocab, (batch, seq), generator=g, dtype=torch.int64).to(dev) for _ in range(2)]

        def timed(steps):
            dist.barrier(); torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for s_ in range(steps):
                eng.step(ids[s_ % 2], ids[s_ % 2])
            e1.record()
            dist.barrier(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item()) / steps

        steps = 5
        for s_ in range(3):
            eng.step(ids[s_ % 2], ids[s_ % 2])
        ms = timed(steps)
        sent = sync.bytes_reduced
        ops.PROFILE = {}
        eng.step(ids[0], ids[0])
        torch.cuda.synchronize()
        prof, ops.PROFILE = ops.PROFILE, None
        big = [e for e in prof.get("gemm", []) if e[3][6] == 2]
        gemm_ms = sum(e[0].elapsed_time(e[1]) for e in big)
        gemm_tf = sum(e[2] for e in big) / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None
        # the same steps with the exchange switched off (ranks drift apart numerically; timing only)
        eng.grad_sync = None
        eng.step(ids[0], ids[0])
        ms_local = timed(steps)
        out = {"workload": desc, "ms_per_step": ms, "tokens_per_s": world * batch * seq / (ms * 1e-3),
               "ms_per_step_without_exchange": ms_local, "exposed_exchange_ms": ms - ms_local,
               "gemm_tflops_in_step": gemm_tf, "gemm_frac_of_sustained_peak": gemm_tf / pk["tf_sustained"] if gemm_tf else None,
               "gradient_bytes_per_step": sent // (3 + steps + 0) if sent else 0, "steps": steps, "warmup": 3,
               "attention": eng.attention}
        del eng, ids
        return out
    except Exception as e:   # the headline must not die with an extra
        return {"workload": desc, "error": f"{type(e).__name__}: {e}"[:300]}


def adamw_large_buffer(ops, dev, pk, n=256 * 1024 * 1024, iters=10) -> dict:
    """Fused AdamW on a 256 Mi-parameter flat buffer, bf16 grads + bf16 shadow: 28 B/param."""
    p = torch.zeros(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    g = torch.full((n,), 0.01, dtype=torch.bfloat16, device=dev)
    sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
    for i in range(3):
        ops.adamw_step(p, m, v, g, step=i + 1, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, shadow=sh)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for i in range(iters):
        ops.adamw_step(p, m, v, g, step=i + 4, lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, shadow=sh)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gbs = 28.0 * n / (ms * 1e-3) / 1e9
    return {"kernel": "adamw (256 Mi params, bf16 grad + shadow, 7.5 GB working set >> L2)", "bound": "hbm",
            "achieved": gbs, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": gbs / pk["hbm_gbs"], "ms_per_launch": ms}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dp-parity", action="store_true", help="N > 1: skip the multi-rank parity check")
    ap.add_argument("--no-extra-configs", action="store_true",
                    help="N > 1: skip the short C3 / C5 runs reported under configs_extra")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--nccl-ctas", type=int, default=8,
                    help="full-parameter workloads: CTAs of the overlapped all-reduce (and SMs x2 left free by the "
                         "backward GEMMs); 0 = default communicator, no reservation")
    ap.add_argument("--tail-layers", type=int, default=1,
                    help="full-parameter workloads: buckets of the bottom N layers go out on the full-width communicator")
    ap.add_argument("--attention", default="auto", choices=["auto", "own", "sdpa"],
                    help="own: this repo's tcgen05 attention kernels (d_head 64 and 128), i.e. a step with no library "
                         "kernel in it; sdpa: torch SDPA (cuDNN, library; A/B measurements); auto = own where built")
    ap.add_argument("--overlap-optimizer", action="store_true",
                    help="full-parameter workloads: per-layer AdamW launches on a second stream beside backward instead "
                         "of one launch after it (measured slower under the power cap; A/B measurement)")
    ap.add_argument("--detail", default="", help="write per-shape GEMM timings of the instrumented pass to this JSON file")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
