"""Generates tests/golden/priority_ref.npz by RUNNING the reference's own PriorityBuffer
(/root/reference/hanabi_agents/rlax_dqn/priority_buffer.py:13-52) on top of the reference's own
sum tree (sum_tree_py.cc, compiled unmodified into oracle/_ref by oracle/build.py).

The file only fails to import here because of `import jax` (:3-5, used by nothing on the live
path) and the `@register_pytree_node_class` decorators of dead classes (:75,103,125): a stub
`jax` module whose decorator is the identity is put in sys.modules, the pybind module is
registered as `sum_tree` (what :10 imports) for the lifetime of this script only, `np.int`
(experience_buffer.py:15, removed in numpy 1.24) is shimmed, and OMP_NUM_THREADS=1 makes the
reference tree sequential (with more threads duplicate indices race).

Recorded: a sequence of add_transitions (fill, exact fill, wrap-around, full-capacity batch),
sample_batch (numpy global RNG seeded per call, the seed is stored) and update_priorities
(including duplicate indices), with every leaf, the total, max/min priority, the cursor and the
ring contents after each operation.

The |td| inputs of update_priorities are chosen so that numpy's float32 `**` (platform libm /
SVML, not correctly rounded everywhere) agrees with the correctly rounded power: values where it
does not are redrawn here, so the fixture pins the algorithm, not one libm build.

Run in the build container only (the GPU box has no /root/reference):

    OMP_NUM_THREADS=1 python tests/golden/make_priority_golden.py
"""
import importlib.util
import os
import sys
import types

os.environ["OMP_NUM_THREADS"] = "1"

import numpy as np  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/hanabi_agents/rlax_dqn"
OUT = os.path.join(HERE, "priority_ref.npz")


def load_reference_priority_buffer():
    sys.path.insert(0, ROOT)
    from oracle import ref_pybind_module
    ref_tree = ref_pybind_module()
    assert ref_tree is not None, "oracle/_ref pybind module missing: python -m oracle.build"
    if not hasattr(np, "int"):
        np.int = int
    jax = types.ModuleType("jax")
    jnp = types.ModuleType("jax.numpy")
    tu = types.ModuleType("jax.tree_util")
    tu.tree_flatten = tu.tree_unflatten = None
    tu.register_pytree_node_class = lambda c: c
    jax.numpy, jax.tree_util = jnp, tu
    saved = {k: sys.modules.get(k) for k in ("jax", "jax.numpy", "jax.tree_util", "sum_tree")}
    sys.modules.update({"jax": jax, "jax.numpy": jnp, "jax.tree_util": tu, "sum_tree": ref_tree})
    try:
        pkg = types.ModuleType("refpkg")
        pkg.__path__ = [REF]
        sys.modules["refpkg"] = pkg
        for name in ("transition", "experience_buffer", "priority_buffer"):
            spec = importlib.util.spec_from_file_location("refpkg." + name, os.path.join(REF, name + ".py"))
            mod = importlib.util.module_from_spec(spec)
            sys.modules["refpkg." + name] = mod
            spec.loader.exec_module(mod)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return sys.modules["refpkg.priority_buffer"].PriorityBuffer


def exact_pow_ok(td, alpha):
    got = (td + 1e-10) ** alpha                      # priority_buffer.py:49 on a float32 array
    x = td.astype(np.float32) + np.float32(1e-10)
    want = np.power(x.astype(np.float64), np.float64(np.float32(alpha))).astype(np.float32)
    return got.dtype == np.float32 and np.array_equal(got, want)


def main():
    PB = load_reference_priority_buffer()
    rng = np.random.default_rng(20261105)
    L, A, cap, alpha = 23, 11, 13, 0.6
    buf = PB(L, A, 1, cap, alpha)
    for b in (buf._obs_tm1_buf, buf._act_tm1_buf, buf._obs_t_buf, buf._lms_t_buf, buf._rew_t_buf,
              buf._terminal_t_buf):
        b[...] = 0                                   # np.empty in the reference
    tree_cap = buf.sum_tree.get_capacity()
    out = {"L": L, "A": A, "cap": cap, "alpha": alpha, "tree_cap": tree_cap}
    ops = []

    def snapshot(k):
        out["s%d_leaves" % k] = np.array(buf.sum_tree.get_values(list(range(tree_cap))), np.float32)
        out["s%d_total" % k] = np.float32(buf.sum_tree.get_total_val())
        out["s%d_max" % k] = np.float64(buf.max_priority)
        out["s%d_min" % k] = np.float64(buf.min_priority)
        out["s%d_size" % k] = buf.size
        out["s%d_oldest" % k] = buf.oldest_entry
        out["s%d_obs_tm1" % k] = buf._obs_tm1_buf.copy()
        out["s%d_obs_t" % k] = buf._obs_t_buf.copy()
        out["s%d_act" % k] = buf._act_tm1_buf.copy()
        out["s%d_rew" % k] = buf._rew_t_buf.copy()
        out["s%d_lms" % k] = buf._lms_t_buf.copy()
        out["s%d_term" % k] = buf._terminal_t_buf.copy()

    def do_add(k, n):
        o1 = (rng.random((n, L)) < 0.3).astype(np.int8)
        o2 = (rng.random((n, L)) < 0.3).astype(np.int8)
        a = rng.integers(0, A, (n, 1)).astype(np.int8)
        r = rng.integers(-2, 3, (n, 1)).astype(np.float64)
        lm = (rng.random((n, A)) < 0.5).astype(np.int8)
        t = (rng.random((n, 1)) < 0.2)
        buf.add_transitions(o1, a, r, o2, lm, t)
        out.update({"a%d_o1" % k: o1, "a%d_o2" % k: o2, "a%d_a" % k: a, "a%d_r" % k: r,
                    "a%d_lm" % k: lm, "a%d_t" % k: t})

    def do_sample(k, B):
        seed = 1000 + k
        np.random.seed(seed)
        idx, prios, tr = buf.sample_batch(B)
        out["p%d_seed" % k] = seed
        out["p%d_B" % k] = B
        out["p%d_idx" % k] = np.array(idx, np.int64)
        out["p%d_prios" % k] = np.asarray(prios, np.float64)
        for name, v in zip(tr._fields, tr):
            out["p%d_%s" % (k, name)] = v
        return np.array(idx, np.int64)

    def do_update(k, idx):
        while True:
            td = np.abs(rng.normal(0, 2.0, len(idx))).astype(np.float32)
            if exact_pow_ok(td, alpha):
                break
        buf.update_priorities(idx, td)
        out["u%d_idx" % k] = np.array(idx, np.int64)
        out["u%d_td" % k] = td

    plan = [("add", 4), ("sample", 4), ("update", None), ("add", 5), ("sample", 6), ("update", None),
            ("add", 4),                       # exact fill: 13 rows
            ("sample", 8), ("update", None),
            ("add", 3),                       # wraps from 0
            ("update", [0, 0, 5, 12, 5, 0]),  # duplicate indices: last one wins, sums in batch order
            ("sample", 13), ("update", None),
            ("add", 12),                      # wrap split 10 + 2 ... new rows take the running max
            ("sample", 32), ("update", None),
            ("add", 13),                      # a full-capacity batch
            ("sample", 7)]
    last_idx = None
    for k, (op, arg) in enumerate(plan):
        ops.append(op)
        if op == "add":
            do_add(k, arg)
        elif op == "sample":
            last_idx = do_sample(k, arg)
        else:
            do_update(k, last_idx if arg is None else np.array(arg, np.int64))
        snapshot(k)
    out["ops"] = np.array(ops)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, "ops:", len(ops))


if __name__ == "__main__":
    main()
