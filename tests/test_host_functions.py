"""Host-side logic of viabel_b200/functions.py (no GPU): the iteration windows handed to vb_rhat_windows must select
exactly the rows the reference's slicing selects (functions.py:12-21, 44-65), including its error behaviour."""
import numpy as np
import pytest

from viabel_b200 import functions as fn


@pytest.mark.parametrize('n_iters', [2, 10, 101, 400])
@pytest.mark.parametrize('warmup', [0, 1, 3, 50, 500, -4, -1000])
def test_kept_window_is_the_references_slice(n_iters, warmup):
    rows = np.arange(n_iters)[warmup:]
    if len(rows) % 2 == 1:
        with pytest.raises(ValueError):
            fn._kept_window(warmup, n_iters)
        return
    begin, kept = fn._kept_window(warmup, n_iters)
    assert kept == len(rows)
    if kept:
        assert begin == rows[0]


def test_halfway_windows_follow_the_reference_loop():
    n_iters, interval, start = 1200, 100, 1000
    want = []
    for i in range(n_iters // interval):
        sub = np.arange(n_iters)[:start + (i + 1) * interval]
        kept = sub[len(sub) // 2:]
        want.append((kept[0], len(kept)))
    got = []
    for i in range(n_iters // interval):
        length = min(start + (i + 1) * interval, n_iters)
        got.append(fn._kept_window(length // 2, length))
    assert got == want
