"""The allocation-free cross-entropy entry point, called straight through the C ABI.

Who needs it: a caller that has already handed all device memory to its own pool, or that
runs inside another framework's allocator, cannot produce the workspace the regular entry
point asks for.  The variant keeps that scratch in a fixed array inside the library, so it
must (a) give the same loss and gradient as the regular entry point, (b) keep doing so when
called again over the previous call's leftovers, and (c) refuse batches beyond its fixed size.

Collected last on purpose (file name), so a run that stops at the first failure has already
exercised every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _both(L, z, targets, vocab, grad_scale):
    """Run the regular (caller workspace) and the allocation-free entry points on copies of z."""
    from b200ft._lib import check
    rows, ld = z.shape
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    za, zb = z.clone(), z.clone()
    la = torch.full((1,), -1.0, device=DEV)
    lb = torch.full((1,), -2.0, device=DEV)
    ws = torch.empty(L.b200ft_softmax_xent_workspace_bytes(rows), dtype=torch.uint8, device=DEV)
    check(L.b200ft_softmax_xent_fwd_bwd(_vp(za), rows, vocab, ld, _vp(targets), -100, grad_scale, _vp(la), _vp(ws),
                                        ws.numel(), stream), "regular entry point")
    check(L.b200ft_softmax_xent_fwd_bwd_noalloc(_vp(zb), rows, vocab, ld, _vp(targets), -100, grad_scale, _vp(lb),
                                                stream), "allocation-free entry point")
    torch.cuda.synchronize()
    return (za, la), (zb, lb)


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    from b200ft import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(11)
    # second call is smaller than the first and has ignored rows: stale per-row slots, a stale
    # count and a used ticket are all left behind by call one
    cases = [(96, 1000, 1024, 1.0), (40, 503, 512, 0.5)]
    for i, (rows, vocab, ld, gs) in enumerate(cases):
        z = (torch.randn(rows, ld, generator=g) * 3).to(torch.bfloat16).to(DEV)
        targets = torch.randint(0, vocab, (rows,), generator=g)
        if i == 1:
            targets[::3] = -100
        (za, la), (zb, lb) = _both(L, z, targets.to(DEV), vocab, gs)
        assert abs(float(la) - float(lb)) <= 1e-5 * max(1.0, abs(float(la))), (i, float(la), float(lb))
        torch.testing.assert_close(zb.float(), za.float(), rtol=2 ** -7, atol=1e-7)
        assert float(lb) > 0.0


def test_noalloc_matches_the_oracle():
    """Not only equal to the regular entry point: equal to the CPU oracle's written-out cross-entropy
    (oracle/gpt2_lora_ref.py softmax_xent_ref), loss to 1e-5 and gradient to one bf16 ulp, on two calls in a row."""
    from b200ft import _lib
    from oracle import gpt2_lora_ref as ref
    L = _lib.lib()
    g = torch.Generator().manual_seed(5)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for rows, vocab, ld in ((33, 50257, 50304), (64, 1000, 1024)):
        z = (torch.randn(rows, vocab, generator=g) * 3).to(torch.bfloat16)
        targets = torch.randint(0, vocab, (rows,), generator=g)
        targets[1] = -100
        loss_ref, dz_ref = ref.softmax_xent_ref(z.float(), targets)
        zz = torch.full((rows, ld), 5.0, dtype=torch.bfloat16)
        zz[:, :vocab] = z
        zd, td, loss = zz.to(DEV), targets.to(DEV), torch.zeros(1, device=DEV)
        rc = L.b200ft_softmax_xent_fwd_bwd_noalloc(_vp(zd), rows, vocab, ld, _vp(td), -100, 1.0, _vp(loss), stream)
        assert rc == 0, L.b200ft_last_error_string()
        assert abs(float(loss) - float(loss_ref)) <= 1e-5 * max(1.0, abs(float(loss_ref)))
        torch.testing.assert_close(zd.cpu().float()[:, :vocab], dz_ref, rtol=2 ** -7, atol=1e-7)
        assert torch.all(zd[:, vocab:] == 0)


@pytest.mark.parametrize("rows", [32768, 65536])
def test_noalloc_at_the_benchmark_row_count_and_at_its_limit(rows):
    """32768 rows is one step of BASELINE.json configs[1] (32 x 1024 tokens); 65536 is the last size the built-in
    scratch holds.  Vocabulary kept small so the buffers stay modest; every row slot of the scratch is in use."""
    from b200ft import _lib
    L = _lib.lib()
    assert rows <= L.b200ft_softmax_xent_noalloc_max_rows()
    g = torch.Generator(device=DEV).manual_seed(rows)
    vocab, ld = 1000, 1024
    z = (torch.randn(rows, ld, generator=g, device=DEV) * 3).to(torch.bfloat16)
    targets = torch.randint(0, vocab, (rows,), generator=g, device=DEV)
    targets[::5] = -100
    (za, la), (zb, lb) = _both(L, z, targets, vocab, 1.0)
    assert abs(float(la) - float(lb)) <= 1e-5 * max(1.0, abs(float(la)))
    assert torch.equal(za, zb)                      # same kernels, same scratch layout: identical bits
    # and against torch's fp32 cross-entropy on the GPU (the CPU oracle would need minutes at this size)
    want = torch.nn.functional.cross_entropy(z[:, :vocab].float(), targets, ignore_index=-100)
    assert abs(float(lb) - float(want)) <= 1e-5 * float(want)


def test_noalloc_refuses_oversize_and_launches_nothing():
    from b200ft import _lib
    L = _lib.lib()
    limit = L.b200ft_softmax_xent_noalloc_max_rows()
    assert limit == 65536
    z = torch.ones(8, 64, dtype=torch.bfloat16, device=DEV)
    t = torch.zeros(8, dtype=torch.int64, device=DEV)
    loss = torch.full((1,), 123.0, device=DEV)
    rc = L.b200ft_softmax_xent_fwd_bwd_noalloc(_vp(z), limit + 1, 64, 64, _vp(t), -100, 1.0, _vp(loss), None)
    assert rc == -3
    assert b"exceeds the built-in scratch" in L.b200ft_last_error_string()
    torch.cuda.synchronize()
    assert float(loss) == 123.0 and torch.all(z == 1)
