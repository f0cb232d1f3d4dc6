"""ctypes binding of the in-tree C-ABI library (include/hanabi_b200.h).

There is no CPU fallback: if the library is missing this module raises, and
every compute entry point fails with HB_ERR_CUDA when no device is usable.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_lib", "libhanabi_b200.so")

HB_OK = 0
HB_ERR_INVALID = -1
HB_ERR_CUDA = -2
HB_ERR_INDEX = -3
HB_ERR_STATE = -4

_lib = None

vp = C.c_void_p
i64 = C.c_int64
i32 = C.c_int
f32 = C.c_float
u64 = C.c_uint64

# name -> (restype, argtypes); kept in step with include/hanabi_b200.h
SIGNATURES = {
    "hb_last_error": (C.c_char_p, []),
    "hb_version": (i32, []),
    "hb_build_arch": (C.c_char_p, []),
    "hb_gate_create": (i32, [i32, C.POINTER(vp)]),
    "hb_gate_destroy": (i32, [vp]),
    "hb_gate_close": (i32, [vp, vp]),
    "hb_gate_open": (i32, [vp]),
    # sum tree
    "hb_st_create": (i32, [i64, i32, C.POINTER(vp)]),
    "hb_st_destroy": (i32, [vp]),
    "hb_st_capacity": (i64, [vp]),
    "hb_st_depth": (i32, [vp]),
    "hb_st_heap": (vp, [vp]),
    "hb_st_update": (i32, [vp, vp, vp, i64, vp]),
    "hb_st_update_range": (i32, [vp, i64, i64, vp, f32, vp]),
    "hb_st_update_range_noalloc": (i32, [vp, i64, i64, vp, f32, vp]),
    "hb_st_get_indices": (i32, [vp, vp, vp, i64, vp]),
    "hb_st_get_values": (i32, [vp, vp, vp, i64, vp]),
    "hb_st_total": (i32, [vp, vp, vp]),
    "hb_st_sample": (i32, [vp, vp, vp, vp, vp, i64, vp]),
    "hb_st_check": (i32, [vp, vp]),
    "hb_st_update_host": (i32, [vp, vp, vp, i64]),
    "hb_st_get_indices_host": (i32, [vp, vp, vp, i64]),
    "hb_st_get_values_host": (i32, [vp, vp, vp, i64]),
    "hb_st_total_host": (i32, [vp, vp]),
    "hb_st_heap_host": (i32, [vp, vp]),
    # environment
    "hb_env_create": (i32, [i32, i32, i32, i32, i32, i32, i64, i32, C.POINTER(vp)]),
    "hb_env_destroy": (i32, [vp]),
    "hb_env_info": (i64, [vp, i32]),
    "hb_env_set_shard": (i32, [vp, i64, i64]),
    "hb_env_set_decks": (i32, [vp, vp, i64, vp]),
    "hb_env_set_decks_host": (i32, [vp, vp, i64]),
    "hb_env_reset": (i32, [vp, vp, vp, vp, vp]),
    "hb_env_step": (i32, [vp, vp, vp, vp, vp, vp, vp]),
    "hb_env_observe": (i32, [vp, i32, vp, vp, vp]),
    "hb_env_obs_len_as": (i64, [vp, i32]),
    "hb_env_observe_as": (i32, [vp, i32, i32, vp, i32, vp]),
    "hb_env_illegal_count": (i32, [vp, C.POINTER(i64), vp]),
    "hb_env_dump": (i32, [vp, vp, vp]),
    "hb_env_unpack_obs": (i32, [vp, vp, vp, i64, vp]),
    "hb_env_unpack_legal": (i32, [vp, vp, vp, i64, vp]),
    "hb_env_step_host": (i32, [vp, vp, vp, vp, vp, vp]),
    "hb_env_step_host_async": (i32, [vp, vp, vp, vp, vp, vp, i32]),
    "hb_rules_act_host_async": (i32, [vp, vp, vp, vp, i32, u64, u64, vp, vp]),
    "hb_env_host_wait": (i32, [vp]),
    "hb_env_step_host_begin": (i32, [vp, vp]),
    "hb_env_step_host_outputs": (i32, [vp, vp, vp, vp, vp]),
    "hb_env_host_wait_actions": (i32, [vp]),
    "hb_env_host_wait_outputs": (i32, [vp, i32]),
    "hb_env_host_slot": (i32, [vp]),
    # rule-based agents
    "hb_rules_act": (i32, [vp, vp, vp, vp, i32, u64, u64, vp, vp, vp, vp]),
    "hb_env_selfplay": (i32, [vp, vp, vp, vp, i32, u64, u64, i32, vp, vp, vp, vp, vp, vp, vp]),
    "hb_st_update_range_dev": (i32, [vp, i64, i64, vp, vp]),
    "hb_st_update_range_indirect": (i32, [vp, vp, i64, vp, vp]),
    # replay shard
    "hb_rb_create": (i32, [i32, i32, i32, i64, i32, f32, i32, C.POINTER(vp)]),
    "hb_rb_destroy": (i32, [vp]),
    "hb_rb_info": (i64, [vp, i32]),
    "hb_rb_tree": (vp, [vp]),
    "hb_rb_add": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, vp]),
    "hb_rb_store_first": (i32, [vp, vp, vp, vp, i64, vp]),
    "hb_rb_add_masked": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, C.POINTER(i64), vp]),
    "hb_rb_add_masked_begin": (i32, [vp, vp, i64, vp]),
    "hb_rb_add_masked_async": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, vp]),
    "hb_rb_add_masked_finish": (i32, [vp, vp, vp, vp, vp, vp, vp, i64, C.POINTER(i64), vp]),
    "hb_rb_gather": (i32, [vp, vp, i64, vp, vp, vp, vp, vp, vp, vp]),
    "hb_rb_check": (i32, [vp, vp]),
    "hb_rb_sample": (i32, [vp, vp, i64, vp, vp, vp, vp]),
    "hb_rb_update_priorities": (i32, [vp, vp, vp, i64, vp]),
    "hb_rb_sample_gather": (i32, [vp, vp, i32, vp, vp, f32, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, vp]),
    "hb_rb_priority_range": (i32, [vp, C.POINTER(f32), C.POINTER(f32), vp]),
    "hb_pack_bits": (i32, [vp, i32, vp, i32, i64, vp]),
    "hb_unpack_bits": (i32, [vp, i32, i32, vp, i64, vp]),
    # q-network
    "hb_gemm_f16": (i32, [vp, i64, vp, i64, i32, i32, i32, i32, f32, vp, i32, vp, i64, vp, i64, vp,
                          i32, f32, f32, f32, vp]),
    "hb_net_create": (i32, [i32, i32, vp, i32, i32, i32, f32, i32, i32, C.POINTER(vp)]),
    "hb_net_destroy": (i32, [vp]),
    "hb_net_info": (i64, [vp, i32]),
    "hb_net_grad": (vp, [vp]),
    "hb_net_set_param": (i32, [vp, i32, i32, vp, i64]),
    "hb_net_get_param": (i32, [vp, i32, i32, vp, i64]),
    "hb_net_get_target_param": (i32, [vp, i32, i32, vp, i64]),
    "hb_net_refresh": (i32, [vp, vp]),
    "hb_net_sync_target": (i32, [vp, vp]),
    "hb_net_set_act_reserve": (i32, [vp, i32]),
    "hb_net_act": (i32, [vp, vp, vp, i32, i64, i32, f32, vp, u64, u64, i64, i32, vp, vp, vp]),
    "hb_net_forward": (i32, [vp, vp, i64, i32, vp, vp, vp]),
    "hb_net_grads": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, f32, f32, i32, vp, vp, vp]),
    "hb_net_adam": (i32, [vp, f32, f32, f32, f32, f32, vp]),
    "hb_net_grads_ex": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, f32, f32, i32, vp, vp, vp, vp]),
    "hb_net_is_weights": (vp, [vp, i32]),
    "hb_net_adam_dyn": (i32, [vp, vp, vp]),
    # fused learner step
    "hb_learner_create": (i32, [vp, vp, i32, i32, C.POINTER(vp)]),
    "hb_learner_destroy": (i32, [vp]),
    "hb_learner_set_graph": (i32, [vp, i32]),
    "hb_learner_info": (i64, [vp, i32]),
    "hb_rb_set_insert_priority": (i32, [vp, f32]),
    "hb_net_grads_rainbow": (i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, f32, vp, vp, vp]),
    "hb_net_adam_tf": (i32, [vp, f32, f32, f32, f32, f32, vp]),
    "hb_net_loss_event": (vp, [vp]),
    "hb_net_kernel_handle": (vp, [i32]),
    "hb_net_learner_buffers": (i32, [vp, i32, vp, vp, vp]),
    "hb_learner_loss": (vp, [vp]),
    "hb_learner_mean_loss": (vp, [vp]),
    "hb_learner_indices": (vp, [vp]),
    "hb_learner_grads": (i32, [vp, vp, vp, f32, f32, i32, vp]),
    "hb_learner_apply": (i32, [vp, f32, f32, f32, f32, f32, vp]),
    "hb_learner_step": (i32, [vp, vp, vp, f32, f32, i32, f32, f32, f32, f32, vp]),
    "hb_learner_set_comm": (i32, [vp, vp, i32]),
    # NVLink all-reduce of the gradient bucket
    "hb_comm_create": (i32, [i32, i32, i32, vp, i64, C.POINTER(vp)]),
    "hb_comm_destroy": (i32, [vp]),
    "hb_comm_export": (i32, [vp, vp]),
    "hb_comm_open": (i32, [vp, vp]),
    "hb_comm_allreduce": (i32, [vp, vp]),
    "hb_comm_info": (i32, [vp, i32]),
    "hb_comm_status": (i32, [vp]),
    "hb_comm_allreduce_range": (i32, [vp, i32, i64, i64, vp]),
    # self-play session
    "hb_session_create": (i32, [i32, i64, i32, C.POINTER(vp)]),
    "hb_session_destroy": (i32, [vp]),
    "hb_session_reset": (i32, [vp, i32, vp]),
    "hb_session_begin_turn": (i32, [vp, i32, vp, vp]),
    "hb_session_end_turn": (i32, [vp, i32, vp, vp, i32, vp]),
    "hb_session_buffer": (vp, [vp, i32, i32]),
    "hb_session_stats": (i32, [vp, vp, vp]),
}


class HanabiB200Error(RuntimeError):
    pass


def lib():
    """Load (once) and return the C-ABI library.  Raises if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HanabiB200Error(
            "CUDA extension %s is missing; build it with "
            "`python -m hanabi_agents_b200.build` (there is no CPU fallback)" % LIB_PATH)
    handle = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(handle, name)
        fn.restype = res
        fn.argtypes = args
    _lib = handle
    return _lib


def last_error():
    return lib().hb_last_error().decode("utf-8", "replace")


def check(rc):
    """Map a C return code to a Python exception."""
    if rc == HB_OK:
        return
    msg = last_error()
    if rc == HB_ERR_INDEX:
        raise IndexError(msg)
    if rc == HB_ERR_INVALID:
        raise ValueError(msg)
    raise HanabiB200Error("hanabi_b200 error %d: %s" % (rc, msg))


_raw_stream = None
_stream_override = []


class use_stream:
    """`with _lib.use_stream(ptr):` -- the library's launches inside go to that cudaStream_t instead of torch's
    current stream.  What `torch.cuda.stream(s)` does for this library's calls at a twentieth of the cost (that
    context manager was 30 us of a 250 us session tick); torch's own allocations and kernels are NOT moved."""

    def __init__(self, stream_ptr):
        self.ptr = int(stream_ptr)

    def __enter__(self):
        _stream_override.append(self.ptr)
        return self

    def __exit__(self, *exc):
        _stream_override.pop()
        return False


def current_stream_ptr():
    """cudaStream_t of torch's current stream (torch is used for hand-off only).  Through torch's raw
    accessor when it has one: `torch.cuda.current_stream()` builds a Stream object and resolves the
    device by name every time, ~5 us a call and ten calls per session tick."""
    global _raw_stream
    if _stream_override:
        return _stream_override[-1]
    import torch
    if _raw_stream is None:
        fn = getattr(torch._C, "_cuda_getCurrentRawStream", None)
        _raw_stream = fn if fn is not None else False
    if _raw_stream:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


class StreamGate:
    """Holds a CUDA stream until `open()` (include/hanabi_b200.h: hb_gate_*): queue the launches of a
    timed region behind it and the region no longer contains the host's launch latency."""

    def __init__(self, device_index):
        self._h = vp()
        check(lib().hb_gate_create(int(device_index), C.byref(self._h)))

    def close(self, stream_ptr=None):
        check(lib().hb_gate_close(self._h, current_stream_ptr() if stream_ptr is None else stream_ptr))

    def open(self):
        check(lib().hb_gate_open(self._h))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value:
            try:
                lib().hb_gate_destroy(h)
            except Exception:
                pass
            self._h = None


class DevView:
    """__cuda_array_interface__ view of library-owned device memory (zero copy into torch)."""

    def __init__(self, ptr, n, typestr, owner):
        self._owner = owner
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


def dev_tensor(ptr, n, typestr, owner, device):
    import torch
    if not ptr:
        raise HanabiB200Error("NULL device buffer")
    return torch.as_tensor(DevView(ptr, n, typestr, owner), device=device)
