"""CPU oracles -- TEST INFRASTRUCTURE, not part of the product.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
`--impl reference` legs may import this package.  Nothing under
hanabi_agents_b200/ does.
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_LIB = os.path.join(HERE, "_build", "liboracle.so")
REF_SEQ_LIB = os.path.join(HERE, "_ref", "libref_sum_tree_seq.so")
REF_OMP_LIB = os.path.join(HERE, "_ref", "libref_sum_tree_omp.so")

vp, i64, i32, f32 = C.c_void_p, C.c_int64, C.c_int, C.c_float
_oracle = None
_refs = {}


def oracle_lib():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_LIB):
            from . import build
            build.build_c_oracles()
        L = C.CDLL(ORACLE_LIB)
        L.so_create.restype = vp
        L.so_create.argtypes = [i64]
        L.so_destroy.argtypes = [vp]
        L.so_capacity.restype = i64
        L.so_capacity.argtypes = [vp]
        L.so_total.restype = f32
        L.so_total.argtypes = [vp]
        L.so_heap.restype = vp
        L.so_heap.argtypes = [vp]
        L.so_update_value.restype = i32
        L.so_update_value.argtypes = [vp, i64, f32]
        L.so_update_values.restype = i32
        L.so_update_values.argtypes = [vp, vp, vp, i64]
        L.so_get_index.restype = i64
        L.so_get_index.argtypes = [vp, f32]
        L.so_get_indices.argtypes = [vp, vp, vp, i64]
        L.so_get_values.restype = i32
        L.so_get_values.argtypes = [vp, vp, vp, i64]
        _oracle = L
    return _oracle


def ref_pybind_module():
    """The reference's own pybind module (sum_tree_py.cc, unmodified, built into oracle/_ref by
    oracle/build.py), loaded from its file WITHOUT touching sys.modules['sum_tree'] -- that top-level
    name belongs to the drop-in shim at the repository root.  None if it was never built."""
    import glob
    import importlib.util
    mods = glob.glob(os.path.join(HERE, "_ref", "sum_tree*.so"))
    if not mods:
        from . import build
        build.build_ref()
        mods = glob.glob(os.path.join(HERE, "_ref", "sum_tree*.so"))
    if not mods:
        return None
    spec = importlib.util.spec_from_file_location("sum_tree", mods[0])
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def ref_lib(omp=False):
    """The reference header compiled into oracle/_ref (None if it was never built)."""
    path = REF_OMP_LIB if omp else REF_SEQ_LIB
    if path not in _refs:
        if not os.path.exists(path):
            from . import build
            build.build_ref()
        if not os.path.exists(path):
            _refs[path] = None
            return None
        L = C.CDLL(path)
        L.ref_st_create.restype = vp
        L.ref_st_create.argtypes = [i64]
        L.ref_st_destroy.argtypes = [vp]
        L.ref_st_capacity.restype = i64
        L.ref_st_capacity.argtypes = [vp]
        L.ref_st_depth.restype = i32
        L.ref_st_depth.argtypes = [vp]
        L.ref_st_total.restype = f32
        L.ref_st_total.argtypes = [vp]
        L.ref_st_update_value.argtypes = [vp, i32, f32]
        L.ref_st_update_values.argtypes = [vp, vp, vp, i64]
        L.ref_st_get_indices.argtypes = [vp, vp, vp, i64]
        L.ref_st_get_values.restype = i32
        L.ref_st_get_values.argtypes = [vp, vp, vp, i64]
        L.ref_st_dump_heap.argtypes = [vp, vp]
        _refs[path] = L
    return _refs[path]


def _i64(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.int64).reshape(-1))


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32).reshape(-1))


class OracleSumTree:
    """C restatement (oracle/sum_tree_oracle.c) with the reference's method names."""

    def __init__(self, capacity):
        self.L = oracle_lib()
        self.h = self.L.so_create(int(capacity))
        if not self.h:
            raise ValueError("bad capacity")

    def __del__(self):
        if getattr(self, "h", None):
            self.L.so_destroy(self.h)
            self.h = None

    def update_value(self, index, value):
        if self.L.so_update_value(self.h, int(index), float(np.float32(value))) != 0:
            raise IndexError(index)

    def update_values(self, indices, values):
        idx, val = _i64(indices), _f32(values)
        if self.L.so_update_values(self.h, idx.ctypes.data, val.ctypes.data, idx.size) != 0:
            raise IndexError("index out of range")

    def get_index(self, q):
        return int(self.L.so_get_index(self.h, float(np.float32(q))))

    def get_indices(self, qs):
        q = _f32(qs)
        out = np.empty(q.size, dtype=np.int64)
        self.L.so_get_indices(self.h, q.ctypes.data, out.ctypes.data, q.size)
        return out.tolist()

    def get_value(self, idx):
        return self.get_values([idx])[0]

    def get_values(self, indices):
        idx = _i64(indices)
        out = np.empty(idx.size, dtype=np.float32)
        if self.L.so_get_values(self.h, idx.ctypes.data, out.ctypes.data, idx.size) != 0:
            raise IndexError("index out of range")
        return out.tolist()

    def get_capacity(self):
        return int(self.L.so_capacity(self.h))

    def get_total_val(self):
        return float(self.L.so_total(self.h))

    def heap(self):
        cap = self.get_capacity()
        buf = (C.c_float * (2 * cap)).from_address(self.L.so_heap(self.h))
        return np.frombuffer(buf, dtype=np.float32).copy()


class RefSumTree:
    """The reference's SumTree<float> itself (oracle/_ref), same method names."""

    def __init__(self, capacity, omp=False):
        self.L = ref_lib(omp)
        if self.L is None:
            raise RuntimeError("oracle/_ref is not built (no /root/reference on this machine)")
        self.h = self.L.ref_st_create(int(capacity))

    def __del__(self):
        if getattr(self, "h", None):
            self.L.ref_st_destroy(self.h)
            self.h = None

    def update_value(self, index, value):
        self.L.ref_st_update_value(self.h, int(index), float(np.float32(value)))

    def update_values(self, indices, values):
        idx, val = _i64(indices), _f32(values)
        self.L.ref_st_update_values(self.h, idx.ctypes.data, val.ctypes.data, idx.size)

    def get_index(self, q):
        return self.get_indices([q])[0]

    def get_indices(self, qs):
        q = _f32(qs)
        out = np.empty(q.size, dtype=np.int64)
        self.L.ref_st_get_indices(self.h, q.ctypes.data, out.ctypes.data, q.size)
        return out.tolist()

    def get_value(self, idx):
        return self.get_values([idx])[0]

    def get_values(self, indices):
        idx = _i64(indices)
        out = np.empty(idx.size, dtype=np.float32)
        if self.L.ref_st_get_values(self.h, idx.ctypes.data, out.ctypes.data, idx.size) != 0:
            raise IndexError("index out of range")
        return out.tolist()

    def get_capacity(self):
        return int(self.L.ref_st_capacity(self.h))

    def get_total_val(self):
        return float(self.L.ref_st_total(self.h))

    def heap(self):
        cap = self.get_capacity()
        out = np.zeros(2 * cap, dtype=np.float32)
        self.L.ref_st_dump_heap(self.h, out.ctypes.data)
        return out
