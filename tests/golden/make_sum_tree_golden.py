"""Generates tests/golden/sum_tree_ref.npz from the REFERENCE sum tree itself.

Run in the build container (needs /root/reference):  python tests/golden/make_sum_tree_golden.py
The reference header is compiled by oracle/build.py into oracle/_ref (no OpenMP =
sequential order) and driven through oracle.RefSumTree.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import RefSumTree  # noqa: E402


def scenario(capacity, seed, n_rounds, batch, dup_frac):
    """fill, then rounds of {stratified sample, update with duplicates}; all inputs recorded."""
    rng = np.random.default_rng(seed)
    t = RefSumTree(capacity)
    cap = t.get_capacity()
    rec = {"capacity": np.int64(capacity), "cap": np.int64(cap)}
    fill_idx = np.arange(cap, dtype=np.int64)
    fill_val = rng.random(cap).astype(np.float32) * np.float32(3.0)
    fill_val[rng.random(cap) < 0.1] = 0.0  # some empty leaves
    t.update_values(fill_idx, fill_val)
    rec["fill_idx"], rec["fill_val"] = fill_idx, fill_val
    rec["heap_after_fill"] = t.heap()
    for r in range(n_rounds):
        keys = np.linspace(1.0 / batch, 1, batch) - rng.uniform(0, 1.0 / batch, batch)
        keys32 = keys.astype(np.float32)
        idx = np.array(t.get_indices(keys32), dtype=np.int64)
        vals = np.array(t.get_values(idx), dtype=np.float32)
        # update: sampled indices (sorted, with natural duplicates) mixed with random ones
        upd_idx = idx.copy()
        m = rng.random(batch) < dup_frac
        upd_idx[m] = rng.integers(0, cap, m.sum())
        if r % 2 == 1:
            rng.shuffle(upd_idx)
        upd_val = (rng.random(batch).astype(np.float32) * np.float32(10.0)) ** np.float32(0.6)
        if r == 2:
            upd_val[::7] = np.float32(1e8)  # large/small mix: leaf != value afterwards
        t.update_values(upd_idx, upd_val)
        rec["keys_%d" % r] = keys32
        rec["idx_%d" % r] = idx
        rec["vals_%d" % r] = vals
        rec["upd_idx_%d" % r] = upd_idx
        rec["upd_val_%d" % r] = upd_val
        rec["total_%d" % r] = np.float32(t.get_total_val())
    rec["heap_final"] = t.heap()
    rec["n_rounds"] = np.int64(n_rounds)
    rec["batch"] = np.int64(batch)
    return rec


if __name__ == "__main__":
    out = {}
    for name, args in {
        "a": (1000, 11, 6, 64, 0.3),    # capacity rounded up 1000 -> 1024
        "b": (4096, 12, 4, 256, 0.1),
        "c": (37, 13, 5, 100, 0.5),     # batch > capacity: many duplicates
    }.items():
        for k, v in scenario(*args).items():
            out["%s_%s" % (name, k)] = v
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "sum_tree_ref.npz")
    np.savez_compressed(path, **out)
    print(path, os.path.getsize(path))
