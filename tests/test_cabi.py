"""CPU: the C-ABI library loads and exports every symbol include/*.h declares."""
import ctypes
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = open(h).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        names.update(re.findall(r"\b(hb_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_exports_every_declared_symbol():
    from hanabi_agents_b200 import _lib
    L = ctypes.CDLL(_lib.LIB_PATH)
    declared = _declared()
    assert len(declared) >= 20
    missing = [n for n in sorted(declared) if not hasattr(L, n)]
    assert not missing, missing
    # and the ctypes signature table covers the header
    assert declared <= set(_lib.SIGNATURES), sorted(declared - set(_lib.SIGNATURES))


def test_version_and_arch_without_gpu():
    from hanabi_agents_b200 import _lib
    assert _lib.lib().hb_version() >= 100
    assert _lib.lib().hb_build_arch() == b"sm_100a"


def test_no_cpu_fallback_without_device():
    import pytest
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    from hanabi_agents_b200 import _lib
    from hanabi_agents_b200.sum_tree import SumTreef
    with pytest.raises(_lib.HanabiB200Error):
        SumTreef(8)


def test_product_never_imports_oracle():
    bad = []
    for dirpath, _, files in os.walk(os.path.join(ROOT, "hanabi_agents_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b|oracle/|liboracle", src, flags=re.M):
                    bad.append(os.path.join(dirpath, f))
    assert not bad, bad


def test_stream_override_nests_and_restores():
    """_lib.use_stream points the library's launches at a given cudaStream_t without touching torch's
    current stream (what the session uses for its per-seat insert streams); host logic only."""
    from hanabi_agents_b200 import _lib
    assert _lib._stream_override == []
    with _lib.use_stream(0x1000):
        assert _lib.current_stream_ptr() == 0x1000
        with _lib.use_stream(0x2000):
            assert _lib.current_stream_ptr() == 0x2000
        assert _lib.current_stream_ptr() == 0x1000
        try:
            with _lib.use_stream(0x3000):
                raise RuntimeError("boom")
        except RuntimeError:
            pass
        assert _lib.current_stream_ptr() == 0x1000
    assert _lib._stream_override == []
