"""dvo_mask_compact_noalloc (include/dvo_cuda.h): the compaction of dvo_mask_compact with its scratch in a static device array and
device buffers of the caller - no counterpart in the reference.  Checked through the C ABI against dvo_mask_compact (integer
work: bit-exact), on calls in a row that share the scratch array.  Collected last on purpose (file name)."""
import numpy as np
import pytest

import dvo_b200
from conftest import make_dvo

pytestmark = pytest.mark.gpu

MAX_N = 1 << 23  # DVO_NOALLOC_MAX_N


def _ctx():
    return make_dvo(width=64, height=48, levels=1, culls=0)


def test_no_alloc_compaction_matches_mask_compact_on_calls_in_a_row():
    rng = np.random.default_rng(77)
    dv = _ctx()
    n_max = 640 * 480 + 77
    d_mask = dv.dev_alloc(n_max); d_idx = dv.dev_alloc(4 * n_max); d_cnt = dv.dev_alloc(4)   # the caller's buffers, made once
    try:
        # a dense mask first, then shorter / sparser ones: the second and third calls find the first call's counts in the scratch
        for n, density in ((n_max, 0.6), (1000, 0.05), (256 * 300, 0.0), (n_max - 1, 0.3), (1, 1.0)):
            mask = (rng.random(n) < density).astype(np.uint8) * rng.integers(1, 256, n).astype(np.uint8)
            want = dv.mask_compact(mask)                       # the existing entry point (host buffers, allocates its scratch)
            dv.copy_to_dev(d_mask, mask)
            dv.copy_to_dev(d_idx, np.full(n, -7, np.int32))
            dv.mask_compact_noalloc(d_mask, n, d_idx, d_cnt)
            dv.sync()
            cnt = np.zeros(1, np.int32); idx = np.empty(n, np.int32)
            dv.copy_to_host(cnt, d_cnt); dv.copy_to_host(idx, d_idx)
            assert int(cnt[0]) == want.size == int(np.count_nonzero(mask))
            assert np.array_equal(idx[:cnt[0]], want)          # integer work: bit-exact
            assert np.all(idx[cnt[0]:] == -7)                  # nothing written past the count
            assert np.array_equal(want, np.flatnonzero(mask).astype(np.int32))
    finally:
        for p in (d_mask, d_idx, d_cnt):
            dv.dev_free(p)
        dv.close()


def test_no_alloc_compaction_two_contexts_and_size_limit():
    dv = _ctx(); dv2 = _ctx()   # the scratch address is resolved per context; both see the same array of the device
    n = 4096 + 5
    d_mask = dv.dev_alloc(n); d_idx = dv.dev_alloc(4 * n); d_cnt = dv.dev_alloc(4)
    try:
        mask = (np.arange(n) % 3 == 0).astype(np.uint8)
        dv.copy_to_dev(d_mask, mask)
        for ctx in (dv, dv2, dv):
            ctx.mask_compact_noalloc(d_mask, n, d_idx, d_cnt)
            ctx.sync()                                         # calls of different contexts must not overlap on the GPU
            cnt = np.zeros(1, np.int32); idx = np.empty(n, np.int32)
            ctx.copy_to_host(cnt, d_cnt); ctx.copy_to_host(idx, d_idx)
            assert np.array_equal(idx[:cnt[0]], np.flatnonzero(mask).astype(np.int32))
        with pytest.raises(dvo_b200.DvoError, match="DVO_NOALLOC_MAX_N"):
            dv.mask_compact_noalloc(d_mask, MAX_N + 1, d_idx, d_cnt)   # rejected before anything is launched
        with pytest.raises(dvo_b200.DvoError):
            dv.mask_compact_noalloc(d_mask, 0, d_idx, d_cnt)
    finally:
        for p in (d_mask, d_idx, d_cnt):
            dv.dev_free(p)
        dv.close(); dv2.close()
