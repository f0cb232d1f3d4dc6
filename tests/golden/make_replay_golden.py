"""Generates tests/golden/replay_ref.npz by running the REFERENCE's own ExperienceBuffer
(/root/reference/hanabi_agents/rlax_dqn/experience_buffer.py, loaded standalone: the package
__init__ imports haiku, and the file needs `np.int`, removed in numpy 1.24, which is shimmed).
Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_replay_golden.py
"""
import importlib.util
import os
import sys
import types

import numpy as np

REF = "/root/reference/hanabi_agents/rlax_dqn"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "replay_ref.npz")


def load_reference_buffer():
    if not hasattr(np, "int"):
        np.int = int
    pkg = types.ModuleType("refpkg")
    pkg.__path__ = [REF]
    sys.modules["refpkg"] = pkg
    for name in ("transition", "experience_buffer"):
        spec = importlib.util.spec_from_file_location("refpkg." + name, os.path.join(REF, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        sys.modules["refpkg." + name] = mod
        spec.loader.exec_module(mod)
    return sys.modules["refpkg.experience_buffer"].ExperienceBuffer


def main():
    EB = load_reference_buffer()
    rng = np.random.default_rng(20260101)
    L, A, cap = 23, 11, 13
    buf = EB(L, A, 1, cap)
    # np.empty in the reference: define the unwritten rows
    for b in (buf._obs_tm1_buf, buf._act_tm1_buf, buf._obs_t_buf, buf._lms_t_buf, buf._rew_t_buf,
              buf._terminal_t_buf):
        b[...] = 0
    out = {"L": L, "A": A, "cap": cap}
    sizes = [4, 5, 4, 1, 12, 13, 3, 7]      # fill, exact fill (13), wraps, a full-capacity batch
    out["sizes"] = np.array(sizes)
    for k, n in enumerate(sizes):
        o1 = (rng.random((n, L)) < 0.3).astype(np.int8)
        o2 = (rng.random((n, L)) < 0.3).astype(np.int8)
        a = rng.integers(0, A, (n, 1)).astype(np.int8)
        r = rng.integers(-2, 3, (n, 1)).astype(np.float64)
        lm = (rng.random((n, A)) < 0.5).astype(np.int8)
        t = (rng.random((n, 1)) < 0.2)
        upd = buf.get_update_indices(n)
        buf.add_transitions(o1, a, r, o2, lm, t)
        out.update({"in%d_o1" % k: o1, "in%d_o2" % k: o2, "in%d_a" % k: a, "in%d_r" % k: r,
                    "in%d_lm" % k: lm, "in%d_t" % k: t, "upd%d" % k: np.array(upd),
                    "st%d_size" % k: buf.size, "st%d_oldest" % k: buf.oldest_entry,
                    "st%d_obs_tm1" % k: buf._obs_tm1_buf.copy(), "st%d_obs_t" % k: buf._obs_t_buf.copy(),
                    "st%d_act" % k: buf._act_tm1_buf.copy(), "st%d_rew" % k: buf._rew_t_buf.copy(),
                    "st%d_lms" % k: buf._lms_t_buf.copy(), "st%d_term" % k: buf._terminal_t_buf.copy()})
    idx = np.array([0, 12, 5, 5, 7, 1])
    tr = buf[idx]
    out["get_idx"] = idx
    for name, v in zip(tr._fields, tr):
        out["get_" + name] = v
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
