"""`sum_tree.SumTreef` with the reference's surface, running on the B200.

Mirrors the pybind class of /root/reference/sum_tree/sum_tree/src/sum_tree_py.cc:7-27
(same nine methods, same argument meaning, `IndexError` from get_value(s) out of
range as `.at()` gives at sum_tree.h:62).  List / numpy arguments go through the
host-buffer C entry points and return Python lists, exactly like the pybind
module; torch CUDA tensors go through the device entry points (stream ordered,
no copy) and return tensors.
"""
import ctypes as C

import numpy as np

from . import _lib


def _is_cuda_tensor(x):
    return hasattr(x, "is_cuda") and x.is_cuda


class SumTreef:
    def __init__(self, capacity, device=None):
        capacity = int(capacity)
        if device is None:
            try:
                import torch
                device = torch.cuda.current_device() if torch.cuda.is_available() else 0
            except ImportError:
                device = 0
        self._device = int(device)
        self._h = C.c_void_p()
        _lib.check(_lib.lib().hb_st_create(capacity, self._device, C.byref(self._h)))

    @classmethod
    def _from_handle(cls, handle, device, owner):
        """Non-owning view of a tree that lives inside another object (the replay shard)."""
        t = cls.__new__(cls)
        t._device = int(device)
        t._h = C.c_void_p(handle)
        t._owner = owner
        return t

    def __del__(self):
        h = getattr(self, "_h", None)
        if getattr(self, "_owner", None) is not None:
            return
        if h is not None and h.value:
            try:
                _lib.lib().hb_st_destroy(h)
            except Exception:
                pass
            self._h = None

    # -- reference surface ------------------------------------------------
    def update_value(self, index, value):
        """updateValue(index, value)  sum_tree.h:32-35"""
        idx = np.array([int(index)], dtype=np.int64)
        val = np.array([value], dtype=np.float32)
        _lib.check(_lib.lib().hb_st_update_host(self._h, idx.ctypes.data, val.ctypes.data, 1))

    def update_values(self, indices, values):
        """updateValues(indices, values)  sum_tree.h:38-44 (sequential order)."""
        if _is_cuda_tensor(indices):
            import torch
            idx = indices.to(torch.int64).contiguous()
            val = values if _is_cuda_tensor(values) else torch.as_tensor(
                np.asarray(values, dtype=np.float32), device=idx.device)
            val = val.to(torch.float32).contiguous()
            if val.numel() != idx.numel():
                raise ValueError("indices and values differ in length")
            _lib.check(_lib.lib().hb_st_update(self._h, idx.data_ptr(), val.data_ptr(),
                                               idx.numel(), _lib.current_stream_ptr()))
            return
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int64).reshape(-1))
        val = np.ascontiguousarray(np.asarray(values, dtype=np.float32).reshape(-1))
        if idx.size != val.size:
            raise ValueError("indices and values differ in length")
        _lib.check(_lib.lib().hb_st_update_host(self._h, idx.ctypes.data, val.ctypes.data, idx.size))

    def update_range(self, start, n, value):
        """Ordered update of leaves [start, start+n) with one value (insert path)."""
        _lib.check(_lib.lib().hb_st_update_range(self._h, int(start), int(n), None, float(value),
                                                 self._stream()))

    def update_range_values(self, start, values):
        """Ordered update of leaves [start, start+len(values)) with one value each (device or host)."""
        import torch
        if not _is_cuda_tensor(values):
            values = torch.as_tensor(np.ascontiguousarray(values, dtype=np.float32),
                                     device=torch.device("cuda", self._device))
        v = values.to(torch.float32).contiguous()
        _lib.check(_lib.lib().hb_st_update_range(self._h, int(start), v.numel(), v.data_ptr(), 0.0,
                                                 self._stream()))

    def get_index(self, quantile):
        """getIndex(quantile)  sum_tree.h:46-50"""
        return self.get_indices([quantile])[0]

    def get_indices(self, quantiles):
        """getIndices(quantiles)  sum_tree.h:52-59"""
        if _is_cuda_tensor(quantiles):
            import torch
            q = quantiles.to(torch.float32).contiguous()
            out = torch.empty(q.numel(), dtype=torch.int64, device=q.device)
            _lib.check(_lib.lib().hb_st_get_indices(self._h, q.data_ptr(), out.data_ptr(),
                                                    q.numel(), _lib.current_stream_ptr()))
            return out
        q = np.ascontiguousarray(np.asarray(quantiles, dtype=np.float32).reshape(-1))
        out = np.empty(q.size, dtype=np.int64)
        _lib.check(_lib.lib().hb_st_get_indices_host(self._h, q.ctypes.data, out.ctypes.data, q.size))
        return out.tolist()

    def get_value(self, idx):
        """getValue(idx)  sum_tree.h:61-63"""
        return self.get_values([idx])[0]

    def get_values(self, indices):
        """getValues(indices)  sum_tree.h:65-72"""
        if _is_cuda_tensor(indices):
            import torch
            idx = indices.to(torch.int64).contiguous()
            out = torch.empty(idx.numel(), dtype=torch.float32, device=idx.device)
            _lib.check(_lib.lib().hb_st_get_values(self._h, idx.data_ptr(), out.data_ptr(),
                                                   idx.numel(), _lib.current_stream_ptr()))
            return out
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int64).reshape(-1))
        out = np.empty(idx.size, dtype=np.float32)
        _lib.check(_lib.lib().hb_st_get_values_host(self._h, idx.ctypes.data, out.ctypes.data, idx.size))
        return out.tolist()

    def get_capacity(self):
        """getCapacity()  sum_tree.h:76 (rounded up to a power of two)."""
        return int(_lib.lib().hb_st_capacity(self._h))

    def get_total_val(self):
        """getTotalVal()  sum_tree.h:74"""
        out = C.c_float()
        _lib.check(_lib.lib().hb_st_total_host(self._h, C.byref(out)))
        return float(out.value)

    def __repr__(self):
        # sum_tree_py.cc:19-22 (std::to_string(float) prints %f)
        return "<SumTree(capacity=%d, maxval=%f)>" % (self.get_capacity(), self.get_total_val())

    # -- device fast path -------------------------------------------------
    def sample(self, quantiles, out=None):
        """get_indices + get_values + get_total_val in one launch (device tensors).
        `out` = (idx int64[n], val float32[n], total float32[1]) reuses caller-owned buffers."""
        import torch
        q = quantiles.to(torch.float32).contiguous()
        if out is not None:
            idx, val, tot = out
        else:
            idx = torch.empty(q.numel(), dtype=torch.int64, device=q.device)
            val = torch.empty(q.numel(), dtype=torch.float32, device=q.device)
            tot = torch.empty(1, dtype=torch.float32, device=q.device)
        _lib.check(_lib.lib().hb_st_sample(self._h, q.data_ptr(), idx.data_ptr(), val.data_ptr(),
                                           tot.data_ptr(), q.numel(), _lib.current_stream_ptr()))
        return idx, val, tot

    def check(self):
        """Raise IndexError if a device-side call saw an out-of-range index."""
        _lib.check(_lib.lib().hb_st_check(self._h, self._stream()))

    def heap(self):
        """Copy of the flat heap (tests): [0] unused, [1] root, leaves at [cap, 2cap)."""
        out = np.empty(2 * self.get_capacity(), dtype=np.float32)
        _lib.check(_lib.lib().hb_st_heap_host(self._h, out.ctypes.data))
        return out

    def _stream(self):
        try:
            return _lib.current_stream_ptr()
        except Exception:
            return None

