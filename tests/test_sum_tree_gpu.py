"""GPU parity: CUDA sum tree (through the C ABI) vs the oracle, bit-exact on indices AND
on every node value; golden vectors generated from the reference header."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(__file__), "golden", "sum_tree_ref.npz")


def _trees(capacity):
    from hanabi_agents_b200.sum_tree import SumTreef
    from oracle import OracleSumTree
    return SumTreef(capacity), OracleSumTree(capacity)


def test_known_answers_reference_gtest():  # test_sum_tree.cc:45-117
    from hanabi_agents_b200.sum_tree import SumTreef
    assert SumTreef(7).get_capacity() == 8 and SumTreef(9).get_capacity() == 16
    t = SumTreef(4)
    t.update_value(0, 1.0)
    assert t.get_value(0) == 1.0 and t.get_total_val() == 1.0
    t.update_value(2, 2.0)
    assert t.get_total_val() == 3.0
    t.update_value(2, 1.0)
    assert t.get_value(2) == 1.0 and t.get_total_val() == 2.0
    t = SumTreef(4)
    t.update_values([0, 1, 2, 3], [1.0, 2.0, 3.0, 4.0])
    assert t.get_total_val() == 10.0
    qs = [0.0, 0.099, 0.1, 0.299, 0.3, 0.599, 0.6, 1.0]
    assert t.get_indices(qs) == [0, 0, 1, 1, 2, 2, 3, 3]
    assert [t.get_index(q) for q in qs] == [0, 0, 1, 1, 2, 2, 3, 3]
    assert repr(t) == "<SumTree(capacity=4, maxval=10.000000)>"
    with pytest.raises(IndexError):
        t.get_value(4)
    with pytest.raises(IndexError):
        t.update_value(4, 1.0)


@pytest.mark.parametrize("name", ["a", "b", "c"])
def test_golden_from_reference(name):
    from hanabi_agents_b200.sum_tree import SumTreef
    g = np.load(GOLD)
    t = SumTreef(int(g[name + "_capacity"]))
    assert t.get_capacity() == int(g[name + "_cap"])
    t.update_values(g[name + "_fill_idx"], g[name + "_fill_val"])
    np.testing.assert_array_equal(t.heap()[1:], g[name + "_heap_after_fill"][1:])
    for r in range(int(g[name + "_n_rounds"])):
        idx = t.get_indices(g["%s_keys_%d" % (name, r)])
        np.testing.assert_array_equal(idx, g["%s_idx_%d" % (name, r)])
        np.testing.assert_array_equal(np.array(t.get_values(idx), np.float32),
                                      g["%s_vals_%d" % (name, r)])
        t.update_values(g["%s_upd_idx_%d" % (name, r)], g["%s_upd_val_%d" % (name, r)])
        assert np.float32(t.get_total_val()) == g["%s_total_%d" % (name, r)]
    np.testing.assert_array_equal(t.heap()[1:], g[name + "_heap_final"][1:])


@pytest.mark.parametrize("capacity", [1, 2, 3, 64, 1000, 1 << 16])
def test_random_updates_every_node(capacity):
    rng = np.random.default_rng(capacity)
    t, o = _trees(capacity)
    cap = o.get_capacity()
    for n in (1, 31, 32, 33, 256, 1024, 1500, 3000):
        idx = rng.integers(0, cap, n)
        if n % 2 == 0:
            idx = np.sort(idx)
        val = (rng.random(n) * 4).astype(np.float32)
        t.update_values(idx, val)
        o.update_values(idx, val)
        np.testing.assert_array_equal(t.heap()[1:], o.heap()[1:])
        q = rng.random(300).astype(np.float32)
        q[0], q[-1] = 0.0, 1.0
        assert t.get_indices(q) == o.get_indices(q)


def test_duplicates_and_cancellation():
    t, o = _trees(8)
    for tree in (t, o):
        tree.update_values([0, 0, 0], [1.0, 2.0, 5.0])
        tree.update_values([1, 1], [1e8, 1.0])
        tree.update_values([3, 2, 3, 2, 3], [0.5, 0.25, 7.0, 1e-3, 0.0])
    np.testing.assert_array_equal(t.heap()[1:], o.heap()[1:])
    assert t.get_value(0) == 5.0 and t.get_value(1) == 0.0


def test_range_update_matches_sequential():
    rng = np.random.default_rng(3)
    t, o = _trees(1 << 14)
    cap = o.get_capacity()
    pre = (rng.random(cap) * 2).astype(np.float32)
    t.update_values(np.arange(cap), pre)
    o.update_values(np.arange(cap), pre)
    for start, n, v in ((0, 1, 0.6), (5, 4097, 0.6), (cap - 700, 700, 1.25), (0, cap, 0.6)):
        t.update_range(start, n, v)
        o.update_values(np.arange(start, start + n), np.full(n, v, np.float32))
        np.testing.assert_array_equal(t.heap()[1:], o.heap()[1:])
    with pytest.raises(IndexError):
        t.update_range(cap - 1, 2, 1.0)


def test_device_tensor_path_and_fused_sample():
    import torch
    rng = np.random.default_rng(9)
    t, o = _trees(1 << 12)
    cap = o.get_capacity()
    val = rng.random(cap).astype(np.float32)
    t.update_values(torch.arange(cap, device="cuda"), torch.from_numpy(val).cuda())
    o.update_values(np.arange(cap), val)
    np.testing.assert_array_equal(t.heap()[1:], o.heap()[1:])
    B = 256
    keys = (np.linspace(1.0 / B, 1, B) - rng.uniform(0, 1.0 / B, B)).astype(np.float32)
    idx, vals, tot = t.sample(torch.from_numpy(keys).cuda())
    torch.cuda.synchronize()
    assert idx.cpu().tolist() == o.get_indices(keys)
    assert vals.cpu().tolist() == o.get_values(o.get_indices(keys))
    assert float(tot.item()) == o.get_total_val()
    assert t.get_indices(torch.from_numpy(keys).cuda()).cpu().tolist() == o.get_indices(keys)
    bad = torch.tensor([0, cap], device="cuda")
    t.get_values(bad)
    with pytest.raises(IndexError):
        t.check()


def test_full_size_properties_cfg3():
    """BASELINE cfg 3 size (capacity 2^22): properties that do not need the oracle at size."""
    import torch
    from hanabi_agents_b200.sum_tree import SumTreef
    cap = 1 << 22
    t = SumTreef(cap)
    rng = np.random.default_rng(0)
    vals = rng.random(cap).astype(np.float32)
    for s in range(0, cap, 65536):
        t.update_values(torch.arange(s, s + 65536, device="cuda"),
                        torch.from_numpy(vals[s:s + 65536]).cuda())
    heap = t.heap()
    np.testing.assert_array_equal(heap[cap:], vals)          # leaves hold what was written
    # every internal node is within fp32 accumulation error of the sum of its children
    lvl = heap[cap // 2:cap]
    np.testing.assert_allclose(lvl, heap[cap:][0::2] + heap[cap:][1::2], rtol=1e-6)
    assert abs(float(heap[1]) - float(vals.astype(np.float64).sum())) / vals.sum() < 1e-4
    B = 256
    keys = (np.linspace(1.0 / B, 1, B) - rng.uniform(0, 1.0 / B, B)).astype(np.float32)
    idx = np.array(t.get_indices(keys))
    assert np.all(np.diff(idx) >= 0) and idx.min() >= 0 and idx.max() < cap   # sortedness
    # prefix-sum property: cumulative mass before the sampled leaf brackets the query
    cs = np.concatenate([[0.0], np.cumsum(vals.astype(np.float64))])
    q = keys.astype(np.float64) * float(heap[1])
    assert np.all(cs[idx] <= q * (1 + 1e-3) + 1.0) and np.all(q <= cs[idx + 1] * (1 + 1e-3) + 1.0)


# ---- the scan formulation of the ordered sums near the root (csrc/sum_tree.cu) -----------------
def _apply_range(t, o, start, vals):
    vals = np.asarray(vals, np.float32)
    t.update_range_values(start, vals)
    o.update_values(np.arange(start, start + len(vals)), vals)
    np.testing.assert_array_equal(t.heap()[1:], o.heap()[1:])


def test_range_scan_fill_from_empty_crosses_every_binade():
    t, o = _trees(1 << 17)
    _apply_range(t, o, 0, np.full(100000, 0.6, np.float32))
    _apply_range(t, o, 100000, np.full((1 << 17) - 100000, 0.6, np.float32))
    _apply_range(t, o, 777, np.full(70001, 0.37, np.float32))       # sums shrink again


def test_range_scan_exact_ties_round_to_even():
    cap = 1 << 17
    t, o = _trees(cap)
    _apply_range(t, o, 0, np.ones(cap, np.float32))                  # root = 2^17 exactly, u = 2^-6
    # diff = 2^-7 = u/2 at the root (a tie at EVERY element), = u one level down, 2u below that ...
    _apply_range(t, o, 0, np.full(cap, 1 + 2.0 ** -7, np.float32))
    rng = np.random.default_rng(5)
    for _ in range(4):   # ties mixed with non-ties and negative diffs, odd starts and lengths
        start = int(rng.integers(0, cap // 2))
        n = int(rng.integers(3000, cap // 2))
        j = rng.integers(-8, 9, n)
        _apply_range(t, o, start, (1 + j * 2.0 ** -9).astype(np.float32))


def test_range_scan_cancellation_large_diffs_and_non_finite():
    cap = 1 << 16
    t, o = _trees(cap)
    rng = np.random.default_rng(6)
    v = rng.random(cap).astype(np.float32)
    v[rng.integers(0, cap, 40)] = 1e8                                # the [PROBE] of SURVEY ST-4
    _apply_range(t, o, 0, v)
    _apply_range(t, o, 0, np.ones(cap, np.float32))                  # huge negative diffs
    v = (rng.random(cap) * 1e-6).astype(np.float32)
    _apply_range(t, o, 0, v)                                         # totals fall through ~20 binades
    v = rng.random(30000).astype(np.float32)
    v[12345] = np.inf
    v[20000] = np.nan
    _apply_range(t, o, 1000, v)
    _apply_range(t, o, 0, rng.random(cap).astype(np.float32))        # NaN stays, as in the reference


@pytest.mark.parametrize("seed", range(6))
def test_range_scan_fuzz(seed):
    rng = np.random.default_rng(100 + seed)
    depth = int(rng.integers(11, 19))
    cap = 1 << depth
    t, o = _trees(cap)
    for _ in range(6):
        n = int(rng.integers(1, cap + 1))
        start = int(rng.integers(0, cap - n + 1))
        kind = int(rng.integers(0, 4))
        if kind == 0:
            v = np.full(n, rng.random() * 2, np.float32)
        elif kind == 1:
            v = (rng.random(n) ** 0.6).astype(np.float32)
        elif kind == 2:
            v = (2.0 ** rng.integers(-12, 6, n)).astype(np.float32)   # powers of two: ties galore
        else:
            v = (rng.random(n) * 10.0 ** rng.integers(-6, 5)).astype(np.float32)
        _apply_range(t, o, start, v)
