"""Data-parallel plumbing: one process per GPU, gradients summed with an all-reduce over
torch.distributed (NCCL over NVLink 5 / NVSwitch on the GPUs; gloo in the CPU tests).

The path shards by batch: every rank holds a full replica and an independent
micro-batch; the only exchange is the all-reduce of the trainable-parameter gradients.  It is
issued per bucket of layers, in reverse layer order, as soon as the bucket's last gradient has
been produced by backward.  ProcessGroupNCCL runs the collective on its own stream, fenced
against the compute stream by an event, so the reduction of bucket k overlaps the backward of
the layers below it; `flush()` makes the compute stream wait before AdamW reads the gradients.
The 1/world averaging is folded into the AdamW kernel's grad_scale.
"""
from __future__ import annotations

import os
import socket

import torch
import torch.distributed as dist


def init_process_group_from_env(backend: str | None = None) -> tuple[int, int, int]:
    """(rank, world, local_rank) from RANK / WORLD_SIZE / LOCAL_RANK (torchrun's contract)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world,
                                    device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local


def free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def layer_buckets(n_layer: int, n_buckets: int) -> list[tuple[int, int]]:
    """Contiguous [lo, hi) layer ranges, highest layers first (they finish backward first)."""
    n_buckets = max(1, min(n_buckets, n_layer))
    edges = [round(i * n_layer / n_buckets) for i in range(n_buckets + 1)]
    return [(edges[i], edges[i + 1]) for i in range(n_buckets - 1, -1, -1)]


def overlap_group(max_ctas: int):
    """A second NCCL communicator over all ranks whose collectives use at most `max_ctas` CTAs.

    Why: the GEMM is a persistent kernel with one CTA (213 KB of shared memory) per SM and a static
    tile schedule.  A default-sized NCCL kernel that lands first takes whole SMs; the GEMM CTAs meant
    for those SMs start only when a neighbour retires, and the launch takes up to twice as long.  With
    the collective confined to a few CTAs and the GEMM grid shrunk by the same SMs (`reserve_ctas` of
    BucketedGradSync) both fit side by side.  Measured on 8 B200, Llama-3-8B full-parameter step
    (16 GB of gradients, one 436 MB bucket per layer; profiles/README.md): default communicator 298.5 ms
    per step with the backward GEMMs at 885 TFLOP/s; 8 CTAs 271.9 ms / 1045 TFLOP/s; 4 CTAs 363 ms (the
    collective, ~12 GB/s per CTA, no longer hides behind backward)."""
    opts = dist.ProcessGroupNCCL.Options()
    opts.config.max_ctas = max_ctas
    opts.config.min_ctas = min(max_ctas, 1)
    return dist.new_group(backend="nccl", pg_options=opts)


class BucketedGradSync:
    """Callable handed to the step engine: `sync(engine, l)` after layer l's gradients are
    complete (l counts down), `sync(engine, -1)` to wait for everything outstanding.
    The engine exposes `grad_views(lo, hi) -> list[Tensor]` of contiguous gradient slices.

    `group` carries the buckets that overlap backward; `tail_group` (default: the same) the bottom
    bucket, which nothing is left to overlap with and which may therefore use a full-width
    communicator.  `reserve_ctas` tells the engine how many SMs to leave to the collective in its
    backward GEMMs (0 = none)."""

    def __init__(self, n_layer: int, n_buckets: int = 2, group=None, tail_group=None, reserve_ctas: int = 0,
                 use_tail_group: bool = False, tail_layers: int = 1):
        self.buckets = layer_buckets(n_layer, n_buckets)
        self.group = group
        self.tail_group = tail_group if use_tail_group else group
        # buckets that start below this layer go out on the tail group: backward has little left to overlap them with,
        # and the CTA-limited communicator by then carries a backlog
        self.tail_layers = max(1, tail_layers)
        self.reserve_ctas = reserve_ctas
        self.pending = []
        self.bytes_reduced = 0

    def __call__(self, engine, l: int) -> None:
        if l < 0:
            self.flush()
            return
        for lo, hi in self.buckets:
            if l == lo:  # backward walks l downward: reaching `lo` completes the bucket
                grp = self.tail_group if lo < self.tail_layers else self.group
                for t in engine.grad_views(lo, hi):
                    self.pending.append(dist.all_reduce(t, op=dist.ReduceOp.SUM, group=grp, async_op=True))
                    self.bytes_reduced += t.numel() * t.element_size()

    def flush(self) -> None:
        for w in self.pending:
            w.wait()  # CUDA: the compute stream waits on the collective's stream; no host block
        self.pending.clear()
