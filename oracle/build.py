"""Build recipe for the CPU oracles (TEST INFRASTRUCTURE, not product).

  oracle/_build/liboracle.so          C restatements (sum tree, Hanabi engine, rules)
  oracle/_ref/libref_sum_tree_seq.so  reference header, no OpenMP (parity oracle)
  oracle/_ref/libref_sum_tree_omp.so  reference header, -fopenmp (CPU baseline)
  oracle/_ref/sum_tree*.so            the reference's own pybind11 module, unmodified

The _ref artefacts are compiled from the sources WHERE THEY LIE under
/root/reference (nothing is copied); they are built only when that tree is
present (this container) and travel to the GPU box as prebuilt files.
"""
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
BUILD = os.path.join(HERE, "_build")
REF_OUT = os.path.join(HERE, "_ref")
REFERENCE = "/root/reference"
REF_INC = os.path.join(REFERENCE, "sum_tree", "sum_tree", "include")
REF_PYBIND_SRC = os.path.join(REFERENCE, "sum_tree", "sum_tree", "src", "sum_tree_py.cc")

C_SOURCES = ["sum_tree_oracle.c", "hanabi_oracle.c", "rules_oracle.c"]


def _run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed: %s\n%s" % (" ".join(cmd), r.stdout + r.stderr))


def _stale(out, srcs):
    if not os.path.exists(out):
        return True
    m = os.path.getmtime(out)
    return any(os.path.getmtime(s) > m for s in srcs)


def build_c_oracles(force=False):
    os.makedirs(BUILD, exist_ok=True)
    out = os.path.join(BUILD, "liboracle.so")
    srcs = [os.path.join(HERE, s) for s in C_SOURCES if os.path.exists(os.path.join(HERE, s))]
    hdrs = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".h")]
    if force or _stale(out, srcs + hdrs):
        gcc = shutil.which("gcc") or "gcc"
        # -ffp-contract=off: never fuse a*b+c, the restatement keeps the reference's rounding
        _run([gcc, "-O2", "-std=c11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off",
              "-o", out] + srcs + ["-lm"])
    return out


def build_ref(force=False):
    """Compile the reference's own sum tree.  Returns dict of built paths (may be empty)."""
    built = {}
    if not os.path.isdir(REF_INC):
        for name in ("libref_sum_tree_seq.so", "libref_sum_tree_omp.so"):
            p = os.path.join(REF_OUT, name)
            if os.path.exists(p):
                built[name] = p
        return built
    os.makedirs(REF_OUT, exist_ok=True)
    gxx = shutil.which("g++") or "g++"
    shim = os.path.join(HERE, "ref_sum_tree_shim.cc")
    hdr = os.path.join(REF_INC, "sum_tree.h")
    seq = os.path.join(REF_OUT, "libref_sum_tree_seq.so")
    omp = os.path.join(REF_OUT, "libref_sum_tree_omp.so")
    base = [gxx, "-O2", "-std=c++14", "-fPIC", "-shared", "-I" + REF_INC, shim]
    if force or _stale(seq, [shim, hdr]):
        _run(base + ["-o", seq])
    if force or _stale(omp, [shim, hdr]):
        _run(base + ["-fopenmp", "-o", omp])
    built["libref_sum_tree_seq.so"] = seq
    built["libref_sum_tree_omp.so"] = omp
    # the reference's own pybind module (sum_tree_py.cc, unmodified), pip pybind11
    try:
        import pybind11
        ext = sysconfig.get_config_var("EXT_SUFFIX")
        mod = os.path.join(REF_OUT, "sum_tree" + ext)
        if force or _stale(mod, [REF_PYBIND_SRC, hdr]):
            _run([gxx, "-O2", "-std=c++14", "-fPIC", "-shared", "-fopenmp",
                  "-I" + pybind11.get_include(), "-I" + sysconfig.get_paths()["include"],
                  "-I" + REF_INC, REF_PYBIND_SRC, "-o", mod])
        built["sum_tree_pybind"] = mod
    except ImportError:
        pass
    return built


if __name__ == "__main__":
    print(build_c_oracles(force="--force" in sys.argv))
    print(build_ref(force="--force" in sys.argv))
