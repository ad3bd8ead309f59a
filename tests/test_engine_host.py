"""Host-side logic of the batched driver that needs no GPU: how a batch is cut into view ranges
(Engine._view_ranges) and when it is not."""
from types import SimpleNamespace

import numpy as np
import pytest

from helpers import mod


def _desc(V, H=840, W=1297):
    return {"H": np.full(V, H), "W": np.full(V, W)}


def test_view_ranges_partition_the_batch_in_order():
    eng = mod("engine")
    opt = eng.RunOptions(aligner="ransac", subsample="adaptive")
    for V in (1, 31, 32, 33, 100, 200, 1001):
        for streams in (1, 2, 3, 4, 8, 64):
            opt.streams = streams
            r = eng.Engine._view_ranges(None, opt, V, _desc(V), True, True)
            assert r[0][0] == 0 and r[-1][1] == V
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))           # contiguous, in order
            assert len(r) <= max(1, streams)
            if len(r) > 1:
                assert min(b - a for a, b in r) >= eng.MIN_VIEWS_PER_RANGE
                assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1   # balanced
            if streams == 1 or V < 2 * eng.MIN_VIEWS_PER_RANGE:
                assert r == [(0, V)]


@pytest.mark.parametrize("change", ["stop_fit", "no_projection", "grad", "defer", "exact", "given_fit", "stride_zero",
                                    "no_subsample", "huge_view", "bad_range"])
def test_view_ranges_fall_back_to_one_range(change):
    """Everything but the plain full pipeline stays one range on the calling stream (the per-pixel mask
    kernel and the partial pipelines do not take offset view pointers)."""
    eng, L = mod("engine"), mod("_lib")
    opt = eng.RunOptions(aligner="ransac", subsample="adaptive", streams=4)
    V, adaptive, have_proj, desc = 200, True, True, _desc(200)
    if change == "stop_fit":
        opt.stop_after = "fit"
    elif change == "no_projection":
        have_proj = False
    elif change == "grad":
        opt.grad_thresh = 0.1
    elif change == "defer":
        opt.defer_emit = True
    elif change == "exact":
        opt.stats_exact = True
    elif change == "given_fit":
        opt.aligner = None
    elif change == "stride_zero":
        opt.subsample, adaptive = 0, False
    elif change == "no_subsample":
        opt.flags &= ~L.KEEP_SUBSAMPLE
    elif change == "huge_view":
        desc = _desc(200, H=70000)
    elif change == "bad_range":
        opt.factor_min, opt.factor_max = 9, 5
    assert eng.Engine._view_ranges(None, opt, V, desc, adaptive, have_proj) == [(0, V)]
    # and the unchanged options do split
    base = eng.RunOptions(aligner="ransac", subsample="adaptive", streams=4)
    assert len(eng.Engine._view_ranges(None, base, V, _desc(200), True, True)) == 4
    static = eng.RunOptions(aligner="lstsqrs", subsample=7, streams=2)
    assert len(eng.Engine._view_ranges(None, static, V, _desc(200), False, True)) == 2
