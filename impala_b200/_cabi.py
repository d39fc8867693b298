"""ctypes binding of the b200rl C ABI (``include/b200rl.h``).

No fallback: if the CUDA library is missing or a call fails this raises
``RuntimeError`` (which also keeps the reference's ``except RuntimeError``
contract at ``agents/impala/learning.py:15-19`` meaningful).
"""
import ctypes
import os

from . import _build

_lib = None

_vp, _i64, _i32, _u32, _u64, _f, _d = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32,
                                       ctypes.c_uint32, ctypes.c_uint64, ctypes.c_float, ctypes.c_double)

class Field(ctypes.Structure):
    """``b200rl_field_t`` (include/b200rl.h)."""
    _fields_ = [("src", _vp), ("dst", _vp), ("row_bytes", _i64), ("kind", _i32), ("gamma", _f)]


FIELD_COPY, FIELD_DONE_TO_DISCOUNT = 0, 1

# name -> (restype, argtypes); mirrors include/b200rl.h one to one
SIGNATURES = {
    "b200rl_version": (_i32, []),
    "b200rl_last_error": (ctypes.c_char_p, []),
    "b200rl_loss_workspace_bytes": (_i64, []),
    "b200rl_trace_offset_bytes": (_i64, []),
    "b200rl_vtrace_impala_loss_fwdbwd": (_i32, [_vp] * 6 + [_i64] * 6 + [_d] + [_f] * 3 + [_u32] + [_vp] * 5),
    "b200rl_vtrace_impala_loss_fwdbwd_noalloc": (_i32, [_vp] * 6 + [_i64] * 6 + [_d] + [_f] * 3 + [_u32] + [_vp] * 4),
    "b200rl_vtrace": (_i32, [_vp] * 5 + [_i64] * 4 + [_d] + [_f] * 2 + [_vp] * 4),
    "b200rl_vtrace_bwd": (_i32, [_vp] * 7 + [_i64] * 4 + [_d, _f] + [_vp] * 4),
    "b200rl_lambda_returns": (_i32, [_vp] * 3 + [_i64] * 4 + [_d] + [_vp] * 2),
    "b200rl_n_step_returns": (_i32, [_vp] * 3 + [_i64] * 5 + [_d] + [_vp] * 2),
    "b200rl_gae": (_i32, [_vp] * 3 + [_i64] * 6 + [_d] + [_vp] * 2),
    "b200rl_ppo_loss_fwdbwd": (_i32, [_vp] * 5 + [_i64] * 3 + [_f] * 2 + [_u32] + [_vp] * 5),
    "b200rl_ppo_is_loss_fwdbwd": (_i32, [_vp] * 7 + [_i64] * 4 + [_f, _d] + [_f] * 4 + [_u32] + [_vp] * 5),
    "b200rl_sample_uniform": (_i32, [_u64, _u32, _u64, _i64, _i64, _vp, _vp]),
    "b200rl_gather_rows": (_i32, [_vp, _i64, _vp, _i64, _i64, _vp, _vp]),
    "b200rl_ring_append": (_i32, [_vp, _i64, _i64, _i64, _vp, _i64, _vp]),
    "b200rl_ring_append_fields": (_i32, [ctypes.POINTER(Field), _i32, _i64, _i64, _i64, _vp]),
    "b200rl_gather_fields": (_i32, [ctypes.POINTER(Field), _i32, _vp, _i64, _i64, _vp]),
    "b200rl_sample_gather_fields": (_i32, [ctypes.POINTER(Field), _i32, _u64, _u32, _u64, _vp, _i64, _i64, _i64, _vp, _vp]),
    "b200rl_counter_add": (_i32, [_vp, _i64, _vp]),
    "b200rl_sumtree_update": (_i32, [_vp, _i64, _vp, _vp, _i64, _vp]),
    "b200rl_sumtree_sample": (_i32, [_vp, _i64, _i64, _u64, _u32, _u64, _i64, _vp, _vp, _vp]),
    "b200rl_quantile_huber_fwdbwd": (_i32, [_vp] * 3 + [_i64] * 3 + [_f] + [_vp] * 3),
    "b200rl_per_td_loss_fwdbwd": (_i32, [_vp] * 4 + [_i64] * 2 + [_f] + [_vp] * 6),
    "b200rl_act_epilogue": (_i32, [_vp, _i64, _i64, _vp, _i64, _u64, _u32, _u64, _vp, _vp]),
    "b200rl_clip_adam_step": (_i32, [_vp] * 4 + [_i64, _f] + [_d] * 4 + [_i64] + [_vp] * 4),
    "b200rl_peer_allreduce_clip_adam_step": (_i32, [_vp] * 7 + [_i32] * 2 + [_vp] * 2 + [_i64] + [_vp] * 2 + [_i64, _f] +
                                             [_d] * 4 + [_i64] * 2 + [_vp] * 4),
    "b200rl_ipc_export": (_i32, [_vp, _vp, _vp]),
    "b200rl_ipc_import": (_i32, [_vp, _i64, _vp]),
    "b200rl_ipc_close": (_i32, [_vp, _i64]),
    "b200rl_obs_u8_to_f32_nhwc": (_i32, [_vp, _vp, _i64, _i64, _i64, _i64, _f, _vp]),
    "b200rl_maxpool3x3s2_fwd": (_i32, [_vp] * 5 + [_i64] * 4 + [_vp]),
    "b200rl_maxpool3x3s2_bwd": (_i32, [_vp] * 3 + [_i64] * 4 + [_vp]),
    "b200rl_bias_act_fwd": (_i32, [_vp] * 5 + [_i64] * 3 + [_vp]),
    "b200rl_relu_bwd_add": (_i32, [_vp] * 5 + [_i64] * 3 + [_vp]),
    "b200rl_conv3x3_wgrad": (_i32, [_vp] * 3 + [_i64] * 6 + [_vp]),
    "b200rl_conv3x3_fwd": (_i32, [_vp] * 3 + [_i64] * 5 + [_i32, _vp]),
    "b200rl_conv3x3_fwd_fused": (_i32, [_vp] * 6 + [_i64] * 5 + [_vp]),
}


def lib_path():
    return _build.LIB_PATH


def load():
    """Load the library (never builds, never falls back)."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise RuntimeError(
            "b200rl CUDA library not built: %s is missing. Run `python -c 'import __graft_entry__ as g; "
            "g.build()'` (needs nvcc). There is no CPU fallback." % path)
    lib = ctypes.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    return load().b200rl_last_error().decode("utf-8", "replace")


def call(name, *args):
    """Call an entry point; non-zero status -> RuntimeError with the library's message."""
    rc = getattr(load(), name)(*args)
    if rc != 0:
        raise RuntimeError("%s failed (status %d): %s" % (name, rc, last_error()))


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()
