"""Build recipe for the in-tree C-ABI library (nvcc, sm_100a only).

`python -m hanabi_agents_b200.build` compiles every translation unit under
csrc/ into `_lib/libhanabi_b200.so`.  The library is built IN-TREE so that it
travels to the GPU box with the repository snapshot; it is git-ignored.
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT_DIR = os.path.join(HERE, "_lib")
LIB_PATH = os.path.join(OUT_DIR, "libhanabi_b200.so")

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
          "-Xptxas", "-v"]
# per-file extra flags (the sum tree must keep the reference's fp32 rounding)
PER_FILE = {
    "sum_tree.cu": ["-fmad=false"],
}


def _nvcc():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: the CUDA extension cannot be built")
    return nvcc


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime():
    m = 0.0
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            if f.endswith((".cuh", ".h", ".hpp")):
                m = max(m, os.path.getmtime(os.path.join(root, f)))
    return m


def build_library(force=False, verbose=False):
    """Compile csrc/*.cu -> _lib/libhanabi_b200.so.  Returns the library path."""
    os.makedirs(OUT_DIR, exist_ok=True)
    nvcc = _nvcc()
    hdr_m = _headers_mtime()
    objs, jobs = [], []
    for src in _sources():
        sp = os.path.join(CSRC, src)
        op = os.path.join(OUT_DIR, src[:-3] + ".o")
        objs.append(op)
        if (not force and os.path.exists(op)
                and os.path.getmtime(op) >= max(os.path.getmtime(sp), hdr_m)):
            continue
        cmd = [nvcc] + ARCH + COMMON + PER_FILE.get(src, []) + ["-c", sp, "-o", op]
        jobs.append((src, cmd))

    def run(job):
        src, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for src, r in ex.map(run, jobs):
                log = os.path.join(OUT_DIR, src[:-3] + ".ptxas.log")
                with open(log, "w") as f:
                    f.write(r.stdout + r.stderr)
                if verbose:
                    sys.stderr.write(r.stdout + r.stderr)
                if r.returncode != 0:
                    raise RuntimeError("nvcc failed for %s:\n%s" % (src, r.stdout + r.stderr))
    need_link = bool(jobs) or not os.path.exists(LIB_PATH) or any(
        os.path.getmtime(o) > os.path.getmtime(LIB_PATH) for o in objs)
    if need_link:
        cmd = [nvcc] + ARCH + ["-shared", "-cudart", "static", "-o", LIB_PATH] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    return LIB_PATH


if __name__ == "__main__":
    p = build_library(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(p)
