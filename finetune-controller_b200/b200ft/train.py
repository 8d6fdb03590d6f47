"""Container entrypoint of the fine-tune job: ``python -m b200ft.train --k=v ...``.

This is what the plugin's ``run_cmd()`` (finetune-controller_b200/plugin/b200_lora.py) starts
inside the controller's ``pytorch`` container (reference
app/jobs/kubeflow/PyTorchJobDeployer.py:93-119).  Contract with the controller:

  * flags: exactly those ``run_cmd()`` emits, incl. ``--dataset_path`` / ``--checkpoint_path``
    (reference app/models/examples/mnist.py:93-95);
  * stdout: a line containing ``Epoch`` before anything the log stream should show
    (reference app/utils/stream_logger.py:404-418, app/core/config.py:26);
  * ``<checkpoint_path>/metrics.csv`` readable by ``pd.read_csv`` (reference
    app/utils/S3Handler.py:252-286), plus ``adapter.pt`` / ``adapter_config.json`` which match
    the default ``store_asset_patterns`` (reference app/models/base/finetuning.py:94-97);
  * exit code 0 on success only — the controller appends ``&& touch done.txt``
    (reference PyTorchJobDeployer.py:30-32); any CUDA/NCCL error propagates as non-zero so the
    operator's restart policy applies.

The pod gets ``--gpus`` devices; the operator's RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT describe PODS
(Master + Workers, reference PyTorchJobDeployer.py:186-249), so this process spawns one worker per local GPU
itself — global rank = pod rank x gpus + local GPU, world = pods x gpus — unless it is already running under
torchrun.  Only global rank 0 (on the Master pod, the one with the s3-sync sidecar) writes artifacts.
"""
from __future__ import annotations

import argparse
import csv
import json
import os
import sys
import time


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="b200ft.train")
    ap.add_argument("--model", default="gpt2-124m")
    ap.add_argument("--mode", default="lora", choices=["lora", "full"],
                    help="lora: adapters on a frozen base; full: every tensor trainable (Llama family)")
    ap.add_argument("--attention", default="auto", choices=["auto", "own", "sdpa"],
                    help="own: this library's tcgen05 attention (d_head 64 / 128 models); sdpa: torch SDPA / cuDNN; "
                         "auto: own wherever it is built")
    ap.add_argument("--lora_r", type=int, default=8)
    ap.add_argument("--lora_alpha", type=float, default=16.0)
    ap.add_argument("--seq_len", type=int, default=1024)
    ap.add_argument("--batch_size", type=int, default=32)
    ap.add_argument("--epochs", type=int, default=1)
    ap.add_argument("--steps_per_epoch", type=int, default=100)
    ap.add_argument("--lr", type=float, default=1e-4)
    ap.add_argument("--weight_decay", type=float, default=0.01)
    ap.add_argument("--beta1", type=float, default=0.9)
    ap.add_argument("--beta2", type=float, default=0.999)
    ap.add_argument("--eps", type=float, default=1e-8)
    ap.add_argument("--max_grad_norm", type=float, default=0.0)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--log_interval", type=int, default=10)
    ap.add_argument("--save_model", action="store_true")
    ap.add_argument("--synthetic", action="store_true")
    ap.add_argument("--save_interval", type=int, default=0,
                    help="steps between resume points (resume.state); 0 = only at the end of each epoch, -1 = never")
    ap.add_argument("--base_weights", default="",
                    help="pretrained base weights (HF *.safetensors file / directory or a .pt state dict); a relative "
                         "path is taken under --dataset_path; empty = random init")
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--nodes", type=int, default=1,
                    help="pods of the PyTorchJob (the plugin's cluster_nodes); > 1 needs the operator's "
                         "RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT in the environment")
    ap.add_argument("--dataset_path", default="/data/dataset")
    ap.add_argument("--checkpoint_path", default="/data/artifacts")
    return ap


def load_token_stream(dataset_path: str):
    """Pre-tokenised data from the dataset mount (the controller's init container copies the
    uploaded object into it, reference PyTorchJobDeployer.py:70-91).  Accepted, first match wins:
    tokens.pt (1-D integer tensor), tokens.bin (little-endian uint16 ids, the nanoGPT/llm.c
    convention), any *.jsonl whose lines carry {"input_ids": [...]}.  None if nothing is there."""
    import glob
    import numpy as np
    import torch
    pt, bn = os.path.join(dataset_path, "tokens.pt"), os.path.join(dataset_path, "tokens.bin")
    if os.path.exists(pt):
        return torch.load(pt).to(torch.int64).flatten()
    if os.path.exists(bn):
        raw = np.fromfile(bn, dtype="<u2")
        return torch.from_numpy(raw.astype(np.int64))
    ids = []
    for path in sorted(glob.glob(os.path.join(dataset_path, "*.jsonl"))):
        with open(path) as f:
            for line in f:
                line = line.strip()
                if line:
                    ids.extend(json.loads(line)["input_ids"])
    return torch.tensor(ids, dtype=torch.int64) if ids else None


def check_token_range(stream, vocab: int) -> None:
    """A token id outside [0, vocab) is a data error (wrong tokenizer for the selected model): the job
    stops with a message instead of training on clamped ids."""
    if stream.numel() == 0:
        raise ValueError("dataset holds no tokens")
    lo, hi = int(stream.min()), int(stream.max())
    if lo < 0 or hi >= vocab:
        raise ValueError(f"dataset token ids span [{lo}, {hi}] but the selected model's vocabulary is [0, {vocab}): "
                         "was it tokenised for another model?")


def batch_from_stream(stream, global_step: int, rank: int, world: int, batch: int, seq: int, vocab: int):
    """Deterministic [batch, seq] window of the token stream for (step, rank): consecutive
    batch*seq spans, ranks interleaved, wrapping around at the end of the stream.  The stream has
    been range-checked once at load (check_token_range)."""
    import torch
    span = batch * seq
    n_chunks = max(1, stream.numel() // span)
    off = ((global_step * world + rank) % n_chunks) * span
    chunk = stream[off: off + span]
    if chunk.numel() < span:
        reps = (span + stream.numel() - 1) // stream.numel()
        chunk = torch.cat([chunk] + [stream] * reps)[:span]
    return chunk.view(batch, seq)


# The resume point must NOT match the controller's artifact patterns (``*.pt`` is one of them, reference
# app/models/base/finetuning.py:94-97): the sidecar re-syncs every match to S3 each AWS_JOB_SYNC_INTERVAL
# (reference PyTorchJobDeployer.py:122-139) and in full-parameter mode this file is 96 GB of optimizer state.
STATE_FILE = "resume.state"


def save_state(eng, checkpoint_path: str, epoch: int, it: int, gstep: int) -> None:
    """Resume point (rank 0).  One file per state tensor, copied to the host and written one at a time — the host never
    holds more than one of them (32 GB each for Llama-3-8B in full-parameter mode, against the plugin's 128 Gi limit) —
    then the small index file last, by rename, so a kill never leaves a resume point that points at half-written data."""
    import torch
    meta, tensors = eng.state_parts()
    tag = f"{gstep:09d}"
    for name, t in tensors.items():
        tmp = os.path.join(checkpoint_path, f"{STATE_FILE}.{name}.{tag}")
        torch.save(t.detach().cpu(), tmp)
    meta = dict(meta, epoch=epoch, it=it, gstep=gstep, tensors=sorted(tensors), tag=tag)
    tmp = os.path.join(checkpoint_path, STATE_FILE + ".tmp")
    torch.save(meta, tmp)
    os.replace(tmp, os.path.join(checkpoint_path, STATE_FILE))
    for f in os.listdir(checkpoint_path):          # tensors of older resume points
        if f.startswith(STATE_FILE + ".") and not f.endswith("." + tag) and not f.endswith(".tmp"):
            os.remove(os.path.join(checkpoint_path, f))


def load_state(checkpoint_path: str):
    """The resume point as one dict (index + tensors), or None.  Tensors are memory-mapped, so loading them into the
    engine streams from disk instead of first materialising 96 GB on the host."""
    import torch
    path = os.path.join(checkpoint_path, STATE_FILE)
    if not os.path.exists(path):
        return None
    st = torch.load(path)
    for name in st.get("tensors", []):
        st[name] = torch.load(os.path.join(checkpoint_path, f"{STATE_FILE}.{name}.{st['tag']}"), mmap=True)
    return st


def worker(args, rank: int, world: int, local: int) -> None:
    import torch
    from . import ddp, gpt2, params

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    lora = gpt2.LoRASpec(args.lora_r, args.lora_alpha)
    opt = gpt2.OptSpec(args.lr, args.beta1, args.beta2, args.eps, args.weight_decay, args.max_grad_norm)
    n_layer = params.MODELS[args.model].n_layer
    full = args.mode == "full"
    # full-parameter gradients are hundreds of MB per layer: one all-reduce bucket per layer
    sync = None
    if world > 1 and full:
        sync = ddp.BucketedGradSync(n_layer, n_buckets=n_layer, group=ddp.overlap_group(8), reserve_ctas=8,
                                    use_tail_group=True)
    elif world > 1:
        sync = ddp.BucketedGradSync(n_layer, n_buckets=1)   # a few MB: one all-reduce of the flat gradient buffer
    # random init, identical on every rank (same seed): the image has no network for checkpoints
    bw = args.base_weights
    if bw and not os.path.isabs(bw):
        bw = os.path.join(args.dataset_path, bw)
    eng, dims = params.build_engine(args.model, lora, opt, batch=args.batch_size, seq=args.seq_len, device=dev,
                                    seed=args.seed, world_size=world, grad_sync=sync, mode=args.mode,
                                    attention=args.attention, base_weights=bw)
    stream = None if args.synthetic else load_token_stream(args.dataset_path)
    if stream is not None:
        check_token_range(stream, dims.vocab)
    B, T = args.batch_size, args.seq_len
    ids_host = torch.empty(B, T, dtype=torch.int64).pin_memory()

    def next_batch(global_step: int):
        if stream is None:
            g = torch.Generator().manual_seed(1234 + args.seed + rank + 7919 * global_step)  # resumable
            ids_host.copy_(torch.randint(0, dims.vocab, (B, T), generator=g, dtype=torch.int64))
        else:
            ids_host.copy_(batch_from_stream(stream, global_step, rank, world, B, T, dims.vocab))
        return ids_host

    # resume: the artifacts volume outlives container restarts inside the pod (restartPolicy
    # OnFailure, backoffLimit 2: reference PyTorchJobDeployer.py:179-189,196)
    start_epoch, start_it, gstep0 = 1, 1, 0
    st = load_state(args.checkpoint_path)
    if st is not None:
        eng.load_state_dict(st)
        start_epoch, start_it, gstep0 = st["epoch"], st["it"] + 1, st["gstep"]
        if start_it > args.steps_per_epoch:
            start_epoch, start_it = start_epoch + 1, 1

    os.makedirs(args.checkpoint_path, exist_ok=True)
    metrics_path = os.path.join(args.checkpoint_path, "metrics.csv")
    writer = fh = None
    if rank == 0:
        resumed = st is not None and os.path.exists(metrics_path)
        fh = open(metrics_path, "a" if resumed else "w", newline="")
        writer = csv.writer(fh)
        if not resumed:
            writer.writerow(["epoch", "step", "loss", "lr", "tokens_per_sec", "elapsed_s"])
        print(f"Epoch {start_epoch - 1}/{args.epochs} starting: model={args.model} mode={args.mode} "
              f"lora_r={args.lora_r} world={world} "
              f"batch/gpu={B} seq={T}" + (f" (resumed at step {gstep0})" if st is not None else ""), flush=True)
    t_start = time.time()
    gstep = gstep0
    for epoch in range(start_epoch, args.epochs + 1):
        t0, tok0 = time.time(), 0
        loss = float("nan")
        for it in range(start_it if epoch == start_epoch else 1, args.steps_per_epoch + 1):
            batch = next_batch(gstep)
            loss = eng.step_host(batch, batch)
            gstep += 1
            tok0 += B * T * world
            if rank == 0 and (it % args.log_interval == 0 or it == args.steps_per_epoch):
                tps = tok0 / max(time.time() - t0, 1e-9)
                print(f"Epoch {epoch}/{args.epochs} step {it}/{args.steps_per_epoch} loss {loss:.4f} "
                      f"tokens/s {tps:.0f}", flush=True)
                writer.writerow([epoch, gstep, f"{loss:.6f}", args.lr, f"{tps:.1f}", f"{time.time() - t_start:.2f}"])
                fh.flush()
            if loss != loss:
                raise RuntimeError("loss is NaN")
            if rank == 0 and args.save_interval >= 0 and (
                    (args.save_interval > 0 and it % args.save_interval == 0) or it == args.steps_per_epoch):
                save_state(eng, args.checkpoint_path, epoch, it, gstep)
    if rank == 0:
        fh.close()
        if args.save_model and full:
            # bf16 weights in the oracle's tensor naming (HF names minus the "model.layers." prefix)
            torch.save({k: v.to(torch.bfloat16) for k, v in eng.params_state().items()},
                       os.path.join(args.checkpoint_path, "model.pt"))
            with open(os.path.join(args.checkpoint_path, "config.json"), "w") as f:
                json.dump({"base_model": args.model, "mode": "full", "dtype": "bfloat16", "steps": gstep,
                           **{k: getattr(dims, k) for k in ("n_layer", "d_model", "n_head", "n_kv_head", "d_ff",
                                                            "vocab", "rms_eps", "rope_theta", "tie_embeddings")}},
                          f, indent=1)
        elif args.save_model:
            from . import weights
            ad = eng.adapters_state()
            torch.save(dict(ad), os.path.join(args.checkpoint_path, "adapter.pt"))     # engine / oracle naming
            weights.save_peft_adapter(args.checkpoint_path, ad, model=args.model, llama=params.is_llama(dims),
                                      r=args.lora_r, alpha=args.lora_alpha, steps=gstep)  # PEFT naming
        print(f"Epoch {args.epochs}/{args.epochs} done: {gstep} steps in {time.time() - t_start:.1f}s", flush=True)


def pod_world(args, env) -> tuple[int, int, str, int | None]:
    """(first global rank of this pod, world size, master address, master port or None = pick a free one).
    One pod: everything is local.  Several pods (plugin cluster_nodes > 1): the Kubeflow training operator
    gives every pod RANK = pod index, WORLD_SIZE = pod count and the Master's address / port."""
    if args.nodes <= 1:
        return 0, args.gpus, "127.0.0.1", None
    try:
        pod, pods = int(env["RANK"]), int(env["WORLD_SIZE"])
        addr, port = env["MASTER_ADDR"], int(env["MASTER_PORT"])
    except KeyError as e:
        raise SystemExit(f"b200ft.train: --nodes={args.nodes} but the operator's {e.args[0]} is not set") from None
    if pods != args.nodes or not 0 <= pod < pods:
        raise SystemExit(f"b200ft.train: --nodes={args.nodes} but WORLD_SIZE={pods}, RANK={pod}")
    return pod * args.gpus, pods * args.gpus, addr, port


def _spawned(local: int, args, rank0: int, world: int, addr: str, port: int) -> None:
    os.environ.update(RANK=str(rank0 + local), LOCAL_RANK=str(local), WORLD_SIZE=str(world),
                      MASTER_ADDR=addr, MASTER_PORT=str(port))
    from . import ddp
    rank, world, local = ddp.init_process_group_from_env("nccl")
    try:
        worker(args, rank, world, local)
    finally:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.destroy_process_group()


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    import torch
    if not torch.cuda.is_available():
        print("b200ft.train: no CUDA device visible; this payload has no CPU path", file=sys.stderr)
        return 2
    if "LOCAL_RANK" in os.environ and int(os.environ.get("WORLD_SIZE", "1")) > 1 and args.gpus > 1:
        from . import ddp  # already under torchrun: one process per GPU exists
        rank, world, local = ddp.init_process_group_from_env("nccl")
        worker(args, rank, world, local)
        return 0
    if args.gpus <= 1 and args.nodes <= 1:
        worker(args, 0, 1, 0)
        return 0
    import torch.multiprocessing as mp
    from . import ddp
    rank0, world, addr, port = pod_world(args, os.environ)
    mp.spawn(_spawned, args=(args, rank0, world, addr, port or ddp.free_port()), nprocs=args.gpus, join=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
