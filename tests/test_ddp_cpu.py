"""Host-side logic of the data-parallel path on CPU: two processes over gloo run the bucketed
all-reduce on a stand-in engine; the result must equal the sum of the per-rank gradients and
every bucket must be issued exactly once, highest layers first."""
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "finetune-controller_b200")


class FakeEngine:
    def __init__(self, n_layer, rank):
        g = torch.Generator().manual_seed(100 + rank)
        self.gA = torch.randn(n_layer * 16, 8, generator=g)
        self.gB = torch.randn(n_layer * 16, 24, generator=g)
        self.calls = []

    def grad_views(self, lo, hi):
        self.calls.append((lo, hi))
        return [self.gA[lo * 16: hi * 16], self.gB[lo * 16: hi * 16]]


def _worker(rank, world, port, n_layer, n_buckets, out_dir):
    sys.path.insert(0, PKG)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from b200ft import ddp
    r, w, _ = ddp.init_process_group_from_env("gloo")
    assert (r, w) == (rank, world)
    eng = FakeEngine(n_layer, rank)
    sync = ddp.BucketedGradSync(n_layer, n_buckets)
    for l in range(n_layer - 1, -1, -1):     # backward order
        sync(eng, l)
    sync(eng, -1)
    torch.save({"gA": eng.gA, "gB": eng.gB, "calls": eng.calls, "bytes": sync.bytes_reduced},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


@pytest.mark.parametrize("n_layer,n_buckets", [(12, 2), (5, 3), (2, 4), (12, 1)])
def test_bucketed_allreduce_world2(tmp_path, n_layer, n_buckets):
    sys.path.insert(0, PKG)
    from b200ft import ddp
    world = 2
    mp.spawn(_worker, args=(world, ddp.free_port(), n_layer, n_buckets, str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    expect_A = sum(FakeEngine(n_layer, r).gA for r in range(world))
    expect_B = sum(FakeEngine(n_layer, r).gB for r in range(world))
    for o in outs:
        torch.testing.assert_close(o["gA"], expect_A)
        torch.testing.assert_close(o["gB"], expect_B)
        covered = sorted(o["calls"])
        assert covered[0][0] == 0 and covered[-1][1] == n_layer
        assert all(covered[i][1] == covered[i + 1][0] for i in range(len(covered) - 1))
        assert o["calls"] == sorted(o["calls"], reverse=True), "buckets must go out top layers first"
        assert o["bytes"] == (expect_A.numel() + expect_B.numel()) * 4


def test_layer_buckets_partition():
    sys.path.insert(0, PKG)
    from b200ft.ddp import layer_buckets
    assert layer_buckets(12, 2) == [(6, 12), (0, 6)]
    assert layer_buckets(12, 1) == [(0, 12)]
    assert layer_buckets(3, 8) == [(2, 3), (1, 2), (0, 1)]


class FakeFullEngine:
    """grad_views() of the full-parameter engine in miniature: a layer slice per bucket, the head on the top
    bucket, the embedding table and the norm gains on the bottom one."""

    def __init__(self, n_layer, rank):
        g = torch.Generator().manual_seed(200 + rank)
        self.n_layer = n_layer
        self.G = torch.randn(n_layer * 32, generator=g)
        self.head, self.embed, self.Gn = (torch.randn(n, generator=g) for n in (24, 40, 7))

    def grad_views(self, lo, hi):
        out = [self.G[lo * 32: hi * 32]]
        if hi == self.n_layer:
            out.append(self.head)
        if lo == 0:
            out += [self.embed, self.Gn]
        return out


def _worker_two_groups(rank, world, port, n_layer, out_dir):
    sys.path.insert(0, PKG)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from b200ft import ddp
    ddp.init_process_group_from_env("gloo")
    overlap = dist.new_group(backend="gloo")      # stands in for ddp.overlap_group(k), which is NCCL-only
    eng = FakeFullEngine(n_layer, rank)
    sync = ddp.BucketedGradSync(n_layer, n_buckets=n_layer, group=overlap, reserve_ctas=8, use_tail_group=True)
    assert sync.group is overlap and sync.tail_group is None and sync.reserve_ctas == 8
    for l in range(n_layer - 1, -1, -1):
        sync(eng, l)
    sync(eng, -1)
    torch.save({k: getattr(eng, k) for k in ("G", "head", "embed", "Gn")}, os.path.join(out_dir, f"g{rank}.pt"))
    dist.destroy_process_group()


def test_per_layer_buckets_on_overlap_group_and_tail_on_default_group(tmp_path):
    """The full-parameter configuration: one bucket per layer on the (CTA-limited) overlap communicator, the bottom
    bucket — nothing left to overlap with — on the default one; every gradient is summed exactly once."""
    sys.path.insert(0, PKG)
    from b200ft import ddp
    world, n_layer = 2, 4
    mp.spawn(_worker_two_groups, args=(world, ddp.free_port(), n_layer, str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(tmp_path / f"g{r}.pt") for r in range(world)]
    for k in ("G", "head", "embed", "Gn"):
        expect = sum(getattr(FakeFullEngine(n_layer, r), k) for r in range(world))
        for o in outs:
            torch.testing.assert_close(o[k], expect)
