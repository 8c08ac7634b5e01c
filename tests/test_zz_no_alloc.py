"""
sb_fm_prefix_f64_noalloc: the FM prefix sum for callers that may not allocate device memory
(a real-time audio callback, a host with a fixed memory budget, a process whose allocator owns
the whole GPU).  Its scratch is a fixed array inside the library, shared by every call and
never cleared, so the things worth checking are exactly these: it gives the same result as
sb_fm_prefix_f64 (which takes caller-allocated scratch), it still does so on a second call
that finds the first call's leftovers in the scratch, and it refuses a range that would not
fit instead of overrunning.

Both entry points launch the same kernels in the same order, so the comparison is exact
(bit-identical doubles), which is tighter than the +-1 LSB the PCM tests use.

This file sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It does not depend on them — it passes alone in a fresh process too.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _prefix(entry, lfo, tb, n0, carry0):
    """Run one of the two entry points through the C ABI; returns (prefix sums, carry after)."""
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    count = lfo.numel()
    out = D.empty(count, torch.float64)
    carry = torch.tensor([carry0], dtype=torch.float64, device=D.device())
    if entry == "alloc":
        scratch = D.empty(C.lib.sb_fm_prefix_scratch_bytes(count), torch.uint8)     # caller-owned scratch
        C.call("sb_fm_prefix_f64", D.ptr(lfo), D.ptr(tb.dev()), tb.nsegs, n0, count, D.ptr(carry),
               D.ptr(out), D.ptr(scratch), D.stream())
    else:
        C.call("sb_fm_prefix_f64_noalloc", D.ptr(lfo), D.ptr(tb.dev()), tb.nsegs, n0, count, D.ptr(carry),
               D.ptr(out), D.stream())
    torch.cuda.synchronize()
    return out.cpu().numpy(), float(carry.item())


def test_noalloc_prefix_matches_allocating_entry_point_twice_in_a_row(cuda):
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    tb = D.timebase_cyc(48000)
    gen = torch.Generator(device="cpu").manual_seed(1234)
    # first call: many tiles, so the built-in scratch is left full of large partial sums
    lfo1 = (torch.rand(1_000_003, generator=gen) * 2 - 1).to(D.device())
    want1, wc1 = _prefix("alloc", lfo1, tb, 0, 0.0)
    got1, gc1 = _prefix("noalloc", lfo1, tb, 0, 0.0)
    assert np.array_equal(got1, want1) and gc1 == wc1
    # second call straight after: different data, fewer tiles, a start offset and a carry —
    # it must not see anything of what the first call left in the shared scratch
    lfo2 = (torch.rand(70_001, generator=gen) * 0.1).to(D.device())
    want2, wc2 = _prefix("alloc", lfo2, tb, 1_000_003, wc1)
    got2, gc2 = _prefix("noalloc", lfo2, tb, 1_000_003, gc1)
    assert np.array_equal(got2, want2) and gc2 == wc2
    # and the sums are what they claim to be: an exclusive running sum of lfo * time step
    step = 1.0 / 48000
    ref = wc1 + np.concatenate([[0.0], np.cumsum(lfo2.cpu().numpy().astype(np.float64) * step)[:-1]])
    np.testing.assert_allclose(got2, ref, rtol=0, atol=1e-9)


def test_noalloc_prefix_refuses_more_than_its_builtin_scratch_holds(cuda):
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    tb = D.timebase_cyc(48000)
    limit = C.lib.sb_fm_prefix_noalloc_max_count()
    assert limit == 8192 * 2048
    tiny = D.zeros(8, torch.float32)
    out = D.empty(8, torch.float64)
    carry = D.zeros(1, torch.float64)
    # the size check comes before anything touches memory, so small buffers are safe here
    with pytest.raises(C.SbError, match="exceeds the built-in scratch"):
        C.call("sb_fm_prefix_f64_noalloc", D.ptr(tiny), D.ptr(tb.dev()), tb.nsegs, 0, limit + 1,
               D.ptr(carry), D.ptr(out), D.stream())
    # an empty range is fine and does nothing
    C.call("sb_fm_prefix_f64_noalloc", D.ptr(tiny), D.ptr(tb.dev()), tb.nsegs, 0, 0, D.ptr(carry),
           D.ptr(out), D.stream())


def test_noalloc_prefix_at_exactly_its_limit(cuda):
    """The largest call the built-in scratch allows (16,777,216 samples: every one of its 8192
    tile slots in use) gives the same sums as the allocating entry point."""
    import torch
    from synthesizer_b200 import _cabi as C, _device as D
    tb = D.timebase_cyc(48000)
    limit = C.lib.sb_fm_prefix_noalloc_max_count()
    gen = torch.Generator(device="cpu").manual_seed(99)
    lfo = (torch.rand(limit, generator=gen) * 2 - 1).to(D.device())
    want, wc = _prefix("alloc", lfo, tb, 0, 0.0)
    got, gc = _prefix("noalloc", lfo, tb, 0, 0.0)
    assert np.array_equal(got, want) and gc == wc
