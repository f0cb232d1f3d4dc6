"""hb_st_update_range_noalloc (include/hanabi_b200.h): the allocation-free variant of the sum tree's
range update (no counterpart in the reference).

Called through the C ABI and compared, bit for bit (the tolerance of every sum-tree test), with the
existing entry point hb_st_update_range on two calls in a row -- the second one overlapping the first,
shorter than it and with per-element values, so that stale contents of the shared scratch array would
show.  Collected last on purpose (the file name): the driver runs pytest -x.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_RANGE = 65536          # HB_ST_NOALLOC_MAX_RANGE


def _tree(L, cap, dev):
    h = C.c_void_p()
    from hanabi_agents_b200 import _lib
    _lib.check(L.hb_st_create(cap, dev, C.byref(h)))
    return h


def _heap(L, h):
    from hanabi_agents_b200 import _lib
    cap = int(L.hb_st_capacity(h))
    out = np.empty(2 * cap, np.float32)
    _lib.check(L.hb_st_heap_host(h, out.ctypes.data))
    return out


@pytest.mark.parametrize("cap", [2 ** 17, 5000])
def test_noalloc_range_update_matches_the_allocating_entry_point(cap):
    import torch
    from hanabi_agents_b200 import _lib
    L = _lib.lib()
    dev = torch.cuda.current_device()
    stream = _lib.current_stream_ptr()
    rng = np.random.default_rng(cap)
    ref, new = _tree(L, cap, dev), _tree(L, cap, dev)
    try:
        tcap = int(L.hb_st_capacity(ref))
        n1 = min(MAX_RANGE, tcap - 7)
        n2 = n1 // 3
        v1 = torch.as_tensor(rng.random(n1).astype(np.float32) * 3, device="cuda")
        v2 = torch.as_tensor(rng.random(n2).astype(np.float32), device="cuda")
        calls = [(7, n1, v1, 0.0),               # per-element values, the largest range allowed
                 (100, n2, v2, 0.0),             # a second call right behind it: shorter, overlapping
                 (3, 1000, None, 0.6)]           # and the insert shape: one fill value (max_priority)
        for k, (start, n, val, fill) in enumerate(calls):
            p = None if val is None else val.data_ptr()
            _lib.check(L.hb_st_update_range(ref, start, n, p, fill, stream))
            _lib.check(L.hb_st_update_range_noalloc(new, start, n, p, fill, stream))
            torch.cuda.synchronize()
            np.testing.assert_array_equal(_heap(L, new), _heap(L, ref), err_msg="call %d" % k)
        # and the tree is a sum tree: root == float64 sum of the leaves to fp32 accuracy
        hp = _heap(L, new)
        assert abs(float(hp[1]) - float(hp[tcap:].astype(np.float64).sum())) <= 1e-4 * float(hp[1])
    finally:
        L.hb_st_destroy(ref)
        L.hb_st_destroy(new)


def test_noalloc_rejects_ranges_beyond_the_fixed_scratch():
    import torch
    from hanabi_agents_b200 import _lib
    L = _lib.lib()
    h = _tree(L, 2 ** 18, torch.cuda.current_device())
    try:
        before = _heap(L, h)
        with pytest.raises(ValueError, match="exceed the fixed scratch"):
            _lib.check(L.hb_st_update_range_noalloc(h, 0, MAX_RANGE + 1, None, 1.0, _lib.current_stream_ptr()))
        with pytest.raises(IndexError):
            _lib.check(L.hb_st_update_range_noalloc(h, 2 ** 18 - 10, 11, None, 1.0, _lib.current_stream_ptr()))
        torch.cuda.synchronize()
        np.testing.assert_array_equal(_heap(L, h), before)      # a rejected call changes nothing
        _lib.check(L.hb_st_update_range_noalloc(h, 0, 0, None, 1.0, _lib.current_stream_ptr()))   # n = 0: no-op
    finally:
        L.hb_st_destroy(h)
