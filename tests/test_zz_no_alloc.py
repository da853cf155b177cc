"""GPU: nbrm_scan_lengths_noalloc against the offsets nbrm_nbc_generate returns.

Who needs this entry point: a caller that packs batches itself and is not
allowed to allocate device memory in its request path, for instance a webhook
or reconcile worker confined to a fixed slice of a GPU it shares with users'
notebook pods.  There a cudaMalloc may fail when the neighbours fill the
device, and it synchronises the whole device even when it succeeds.  The
entry point therefore keeps its only scratch (4096 tile sums, 32 KiB) in an
array inside the library and takes every other buffer from the caller.

What is checked: nbrm_nbc_generate lays its objects out with the same scan
(using scratch it allocated), so the lengths of its objects, fed through the
no-allocation entry point, must give back exactly its offsets.  Integers, so
the comparison is bit-exact, like every other parity test here.  Two calls
are made back to back, the second on a smaller batch, so the second call runs
over scratch that still holds the first call's tile sums.

This file sorts last on purpose: with `pytest -x` every other test has already
run if this one fails.  It also passes when run on its own in a fresh process.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _offsets_via_noalloc(pkg, lens):
    import torch
    n = len(lens)
    d_lens = torch.from_numpy(lens.astype(np.uint32).view(np.int32).copy()).cuda()
    d_off = torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream().cuda_stream
    rc, msg = pkg.scan_lengths_noalloc(d_lens.data_ptr() if n else None, n, d_off.data_ptr(), stream)
    assert rc == 0, msg
    torch.cuda.synchronize()
    return d_off.cpu().numpy().view(np.uint64)


def test_noalloc_scan_matches_generate_offsets_twice_in_a_row(ctx, pkg, wl):
    ctx.set_env({"USE_ISTIO": "true"})
    for n_notebooks in (20000, 700):          # 60 000 lengths (59 tiles), then 2 100 (3 tiles)
        nbs = wl.corpus("fuzz", n_notebooks, start=n_notebooks)
        res = ctx.nbc_generate(pkg.Docs([wl.dumps(nb) for nb in nbs]), 7 | 8)
        lens = np.diff(res.off).astype(np.uint32)
        assert len(lens) == 3 * n_notebooks and int(lens.max()) > 0
        got = _offsets_via_noalloc(pkg, lens)
        assert got.dtype == res.off.dtype and np.array_equal(got, res.off)


def test_noalloc_scan_leaves_device_memory_alone(pkg):
    """The claim in the name: the amount of free device memory is the same before and after a
    call (the first call on a device may load the library's module, so one call is made first)."""
    import torch
    lens = np.arange(1, 300001, dtype=np.uint32)
    d_lens = torch.from_numpy(lens.view(np.int32).copy()).cuda()
    d_off = torch.zeros(len(lens) + 1, dtype=torch.int64, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    rc, msg = pkg.scan_lengths_noalloc(d_lens.data_ptr(), len(lens), d_off.data_ptr(), stream)
    assert rc == 0, msg
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        rc, msg = pkg.scan_lengths_noalloc(d_lens.data_ptr(), len(lens), d_off.data_ptr(), stream)
        assert rc == 0, msg
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before
    assert int(d_off[-1].item()) == int(lens.astype(np.uint64).sum())


def test_noalloc_scan_edges_and_limit(pkg):
    import torch
    assert list(_offsets_via_noalloc(pkg, np.zeros(0, dtype=np.uint32))) == [0]
    assert list(_offsets_via_noalloc(pkg, np.array([5], dtype=np.uint32))) == [0, 5]
    big = np.full(1025, 0xffffffff, dtype=np.uint32)       # sums need 64 bits, crosses a tile edge
    assert int(_offsets_via_noalloc(pkg, big)[-1]) == 1025 * 0xffffffff
    # beyond the fixed scratch: refused with a message, nothing launched, outputs untouched
    d_off = torch.full((4,), -1, dtype=torch.int64, device="cuda")
    d_lens = torch.zeros(4, dtype=torch.int32, device="cuda")
    rc, msg = pkg.scan_lengths_noalloc(d_lens.data_ptr(), pkg.SCAN_NOALLOC_MAX_N + 1, d_off.data_ptr(), None)
    torch.cuda.synchronize()
    assert rc == -3 and "4194304" in msg and "at most" in msg
    assert d_off.cpu().tolist() == [-1, -1, -1, -1]
