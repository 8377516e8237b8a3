"""raht_channel_minmax_static (-m gpu; collected last on purpose): the allocation-free variant of raht_channel_minmax,
whose scratch is a file-scope device array of the library.  Called through the C-ABI with caller-owned inputs and
outputs only, twice in a row (the second call finds the first call's contents in the shared scratch), and compared
with the workspace entry point.  Minima and maxima are selections, not arithmetic: the tolerance is zero."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _both(T):
    from raht_b200 import _lib as L
    from raht_b200._lib import lib, check, ptr, stream
    n, D = T.shape
    tag = L.dtype_tag(T.dtype)
    ws_bytes = lib.raht_channel_minmax_workspace_bytes(n, D)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=T.device)
    ref = torch.empty(2, D, dtype=torch.float64, device=T.device)
    got = torch.full((2, D), float("nan"), dtype=torch.float64, device=T.device)
    check(lib.raht_channel_minmax(ptr(T), tag, n, D, ptr(ref), ptr(ws), ws_bytes, stream()), "raht_channel_minmax")
    # Nothing may be allocated during the call: torch's count of allocations ever made stays put and the bytes in use do not
    # grow.  The garbage collector is held off for the window -- after the other tests it may free cyclic garbage at any
    # moment, which moved `memory_allocated` under an earlier form of this check (equality).
    import gc
    torch.cuda.synchronize()
    gc.collect()
    gc.disable()
    try:
        allocs = torch.cuda.memory_stats(T.device).get("allocation.all.allocated", 0)
        before = torch.cuda.memory_allocated(T.device)
        check(lib.raht_channel_minmax_static(ptr(T), tag, n, D, ptr(got), stream()), "raht_channel_minmax_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_stats(T.device).get("allocation.all.allocated", 0) == allocs
        assert torch.cuda.memory_allocated(T.device) <= before
    finally:
        gc.enable()
    return got.cpu().numpy(), ref.cpu().numpy()


def test_static_scratch_minmax_matches_the_workspace_entry_point_on_two_calls_in_a_row():
    from raht_b200._lib import lib
    assert lib.raht_channel_minmax_static_max_channels() == 64
    with torch.cuda.device(DEV):
        g = torch.Generator(device=DEV).manual_seed(77)
        # first call: many CTAs, all 64 channels of the scratch written; second: a smaller, different matrix whose
        # result must not see the leftovers; third: fp32 input, one row
        cases = [torch.randn(300_000, 59, generator=g, device=DEV, dtype=torch.float64) * 1e3,
                 torch.rand(1_237, 7, generator=g, device=DEV, dtype=torch.float64) - 0.25,
                 torch.randn(1, 64, generator=g, device=DEV, dtype=torch.float32)]
        for T in cases:
            got, ref = _both(T)
            assert np.array_equal(got, ref)
            host = T.double().cpu().numpy()
            assert np.array_equal(got[0], host.min(0)) and np.array_equal(got[1], host.max(0))


def test_static_scratch_minmax_rejects_what_does_not_fit():
    from raht_b200 import _lib as L
    from raht_b200._lib import lib, ptr, stream
    with torch.cuda.device(DEV):
        T = torch.zeros(16, 65, dtype=torch.float64, device=DEV)
        out = torch.empty(2, 65, dtype=torch.float64, device=DEV)
        rc = lib.raht_channel_minmax_static(ptr(T), L.dtype_tag(T.dtype), 16, 65, ptr(out), stream())
        assert rc != 0
        assert b"64 channels at most" in lib.raht_last_error()
        assert lib.raht_channel_minmax_static(None, L.dtype_tag(T.dtype), 16, 8, ptr(out), stream()) != 0
        assert lib.raht_channel_minmax_static(ptr(T), L.dtype_tag(T.dtype), 0, 8, ptr(out), stream()) != 0
