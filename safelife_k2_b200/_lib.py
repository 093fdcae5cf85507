"""
ctypes loader for libsafelife_b200.so (the C ABI declared in include/safelife_b200.h).

There is NO CPU fallback: if the CUDA library is missing or a call fails, this raises.
PyTorch is used by the callers only for device memory, streams and torch.distributed.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# SLB_LIB_PATH: another build of the same library (A/B runs of compile-time knobs); never a fallback
LIB_PATH = os.environ.get("SLB_LIB_PATH") or os.path.join(_HERE, "libsafelife_b200.so")

SLB_OK = 0
SLB_ERR_INVALID = -1
SLB_ERR_CUDA = -2
SLB_ERR_UNSUPPORTED = -3
SLB_ERR_OVERFLOW = -4


class SlbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libsafelife_b200 error %d: %s" % (code, msg))
        self.code = code


class Pool(C.Structure):
    """slb_pool (include/safelife_b200.h)"""
    _fields_ = [
        ("n_levels", C.c_int32), ("h", C.c_int32), ("w", C.c_int32), ("words_per_row", C.c_int32),
        ("board", C.c_void_p), ("goals", C.c_void_p),
        ("agent_x", C.c_void_p), ("agent_y", C.c_void_p), ("orientation", C.c_void_p),
        ("spawn_prob", C.c_void_p), ("min_performance", C.c_void_p),
        ("baseline", C.c_void_p), ("exit_mask", C.c_void_p),
        ("board_res", C.c_void_p), ("goals_res", C.c_void_p), ("plane_bits", C.c_void_p),
        ("plane_groups", C.c_uint32 * 2),
        ("level_consts", C.c_void_p), ("external_writes", C.c_uint32), ("reserved", C.c_uint32),
    ]


# every symbol include/safelife_b200.h declares: name -> (restype, argtypes)
_vp, _i, _u64, _sz = C.c_void_p, C.c_int, C.c_uint64, C.c_size_t
_PP = C.POINTER(Pool)
SYMBOLS = {
    "slb_last_error": (C.c_char_p, []),
    "slb_abi_version": (_i, []),
    "slb_launch_count": (_u64, []),
    "slb_pool_prepare": (_i, [_PP, _vp]),
    "slb_resident_bytes": (_sz, [_i, _i]),
    "slb_resident_import": (_i, [_vp, _vp, _i, _i, _i, _PP, _i, _vp]),
    "slb_resident_export": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "slb_advance_board": (_i, [_vp, _vp, _i, _i, _i, _vp, _u64, _u64, _vp, _i, _vp, _vp]),
    "slb_advance_board_host": (_i, [_vp, _vp, _i, _i, _i, _vp, _u64, _u64, _vp, _i]),
    "slb_env_reset": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _vp, _vp]),
    "slb_execute_action": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _vp]),
    "slb_env_set_min_performance": (_i, [_i, _vp, _vp, C.c_double, _vp]),
    "slb_game_advance": (_i, [_PP, _i, _vp, _vp, _vp, _u64, _u64, _vp]),
    "slb_points_and_performance": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "slb_update_exit_colors": (_i, [_PP, _i, _vp, _vp, _vp, _vp]),
    "slb_env_step": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _i, _u64, _u64, _vp, _vp, _i, _vp]),
    "slb_env_rollout": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _i, _i, _u64, _u64, _vp, _vp, _i, _vp]),
    "slb_env_step_host": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _i, _u64, _u64, _vp, _vp, _i, _vp]),
    "slb_env_step_host_async": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _i, _u64, _u64, _vp, _vp, _i, _vp]),
    "slb_stream_wait": (_i, [_vp]),
    "slb_stepper_create": (_i, [_PP, _i, _vp, _vp, _vp, _i, _u64, _u64, _i, _vp, C.POINTER(_vp)]),
    "slb_stepper_step": (_i, [_vp, _vp, _vp, _vp]),
    "slb_stepper_wait": (_i, [_vp]),
    "slb_stepper_destroy": (None, [_vp]),
    "slb_stepper_bind_outputs": (_i, [_vp, _vp, _vp]),
    "slb_side_effect_counts": (_i, [_vp, _vp, _i, _i, _i, _vp, _i, _vp, _u64, _u64, _vp, _i, _vp,
                                    _vp, _vp, _vp, _vp]),
    "slb_accumulate_stats": (_i, [_i, _vp, _vp, _vp, _vp, _vp]),
    "slb_emd_sizes": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _vp]),
    "slb_emd_solve": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _i, C.c_double, _vp, _i, _i, _sz, _vp, _sz,
                           _vp, _vp, _vp, _vp]),
    "slb_emd_solve_static": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _i, C.c_double, _vp, _i, _i, _sz,
                                  _vp, _vp, _vp]),
    "slb_get_obs": (_i, [_PP, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _vp]),
    "slb_env_step_obs": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _i, _u64, _u64, _vp, _vp, _i, _i, _i, _i, _vp, _i, _vp, _vp]),
    "slb_get_obs_channels": (_i, [_PP, _i, _vp, _vp, _vp, _i, _i, _i, _vp, _i, _vp, _vp, _vp]),
    "slb_side_effect_penalty": (_i, [_PP, _i, _vp, _vp, _vp, _vp, _vp]),
    "slb_obs_channels": (_i, [_vp, _sz, _vp, _i, _vp, _vp]),
    "slb_dev_stepper_create": (_i, [_PP, _i, _vp, _vp, _vp, _i, _u64, _u64, _i, _vp, _i, _i, _i, _vp, _i,
                                    _vp, C.POINTER(_vp)]),
    "slb_dev_stepper_step": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "slb_dev_stepper_destroy": (None, [_vp]),
    "slb_microbench_int": (_i, [_i, _i, _i, _vp, C.POINTER(_u64), _vp]),
}

_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built — no fallback."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "libsafelife_b200.so not found at %s — build it with "
                "`python -c 'import __graft_entry__ as g; g.build()'` (nvcc, sm_100a). "
                "There is no CPU fallback." % LIB_PATH)
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        if lib.slb_abi_version() != 5:
            raise ImportError("libsafelife_b200.so ABI version mismatch")
        _lib = lib
    return _lib


def check(rc):
    if rc != SLB_OK:
        raise SlbError(rc, load().slb_last_error().decode("utf-8", "replace"))


def launch_count():
    return int(load().slb_launch_count())


def ptr(t):
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if t is None:
        return None
    if hasattr(t, "data_ptr"):
        return t.data_ptr()
    return t.ctypes.data


def current_stream():
    import torch

    return torch.cuda.current_stream().cuda_stream
