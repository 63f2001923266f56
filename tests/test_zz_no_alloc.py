"""kssd_b200_set_subtract_noalloc (rabbitkssd_b200/csrc/noalloc.cu): the allocation-free variant of set_subtract, through
the C-ABI, against the existing entry point (kssd_b200_set_subtract) -- bit-exact, on two calls in a row (the second one
smaller, so it runs over the first call's leftovers in the shared scratch array), plus the size limits.
Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_HASHES = 4 << 20
MAX_GENOMES = 64 << 10


def _case(rng, sizes, ref_set):
    qs = []
    for n in sizes:
        own = rng.integers(0, 1 << 28, n, dtype=np.uint64).astype(np.uint32)
        take = rng.choice(ref_set, n // 2) if n else np.zeros(0, np.uint32)
        q = np.concatenate([own, take])
        rng.shuffle(q)  # the order of the survivors inside a genome must be kept
        qs.append(q)
    off = np.concatenate([[0], np.cumsum([len(q) for q in qs])]).astype(np.uint64)
    return np.concatenate(qs) if qs else np.zeros(0, np.uint32), off


def _noalloc(ctx, torch, q, off, ref_set):
    dev = torch.device("cuda:0")
    d_q = torch.from_numpy(q.view(np.int32)).to(dev) if len(q) else torch.zeros(1, dtype=torch.int32, device=dev)
    d_off = torch.from_numpy(off.view(np.int64)).to(dev)
    d_ref = torch.from_numpy(ref_set.view(np.int32)).to(dev) if len(ref_set) else torch.zeros(1, dtype=torch.int32, device=dev)
    d_out = torch.full((max(len(q), 1),), -1, dtype=torch.int32, device=dev)
    d_out_off = torch.full((len(off),), -1, dtype=torch.int64, device=dev)
    torch.cuda.synchronize()  # the buffers are the caller's; the library works on its own stream
    ctx.set_subtract_noalloc(d_q.data_ptr(), d_off.data_ptr(), len(off) - 1, len(q), d_ref.data_ptr(), len(ref_set),
                             d_out.data_ptr(), d_out_off.data_ptr())
    ctx.sync()
    goff = d_out_off.cpu().numpy().view(np.uint64)
    return d_out.cpu().numpy().view(np.uint32)[:int(goff[-1])], goff


def test_set_subtract_noalloc_equals_set_subtract():
    import torch
    from rabbitkssd_b200 import _lib, api
    ctx = api.Context(0)
    try:
        rng = np.random.default_rng(2024)
        ref_set = np.unique(rng.integers(0, 1 << 28, 400000, dtype=np.uint64).astype(np.uint32))
        # two calls in a row: a large one, then a smaller one with empty genomes over the first one's leftovers
        for sizes in ([90000, 0, 1221, 250000, 7], [5, 0, 333, 0, 40000, 1, 0]):
            q, off = _case(rng, sizes, ref_set)
            want_h, want_off = ctx.set_subtract(q, off, ref_set)
            got_h, got_off = _noalloc(ctx, torch, q, off, ref_set)
            assert np.array_equal(got_off, want_off)
            assert np.array_equal(got_h, want_h)
            assert 0 < len(got_h) < len(q)  # something was dropped, something survived
        # nothing to subtract, and no genomes at all
        q, off = _case(rng, [100, 0, 50], ref_set)
        got_h, got_off = _noalloc(ctx, torch, q, off, np.zeros(0, np.uint32))
        assert np.array_equal(got_h, q) and np.array_equal(got_off, off)
        got_h, got_off = _noalloc(ctx, torch, np.zeros(0, np.uint32), np.zeros(1, np.uint64), ref_set)
        assert len(got_h) == 0 and list(got_off) == [0]
    finally:
        ctx.close()


def test_set_subtract_noalloc_rejects_sizes_beyond_the_scratch():
    import torch
    from rabbitkssd_b200 import _lib, api
    ctx = api.Context(0)
    try:
        d = torch.zeros(16, dtype=torch.int64, device="cuda:0")
        p = d.data_ptr()
        with pytest.raises(_lib.KssdError, match="exceed the fixed scratch"):
            ctx.set_subtract_noalloc(p, p, 1, MAX_HASHES + 1, p, 0, p, p)
        with pytest.raises(_lib.KssdError, match="exceed the fixed scratch"):
            ctx.set_subtract_noalloc(p, p, MAX_GENOMES + 1, 8, p, 0, p, p)
        with pytest.raises(_lib.KssdError, match="bad arguments"):
            ctx.set_subtract_noalloc(p, 0, 1, 8, p, 0, p, p)
    finally:
        ctx.close()
