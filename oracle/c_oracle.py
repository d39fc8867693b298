"""ctypes binding of the plain-C oracle (``oracle/c/oracle_c.c``) -- TEST ORACLE ONLY."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_c.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "c", "oracle_c.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", os.path.join(_HERE, "c")])
    return _SO


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def philox4x32_10(ctr, key):
    c = np.asarray(ctr, dtype=np.uint32)
    k = np.asarray(key, dtype=np.uint32)
    o = np.zeros(4, dtype=np.uint32)
    lib().oracle_philox4x32_10(_p(c), _p(k), _p(o))
    return tuple(int(x) for x in o)


def sample_uniform(seed, stream_id, draw_offset, n, size):
    out = np.zeros(n, dtype=np.int64)
    rc = lib().oracle_sample_uniform(ctypes.c_uint64(seed), ctypes.c_uint32(stream_id),
                                     ctypes.c_uint64(draw_offset), ctypes.c_int64(n),
                                     ctypes.c_int64(size), _p(out))
    if rc != 0:
        raise RuntimeError("oracle_sample_uniform rc=%d" % rc)
    return out


def gather_rows(src, idx):
    src = np.ascontiguousarray(src)
    idx = np.ascontiguousarray(idx, dtype=np.int64)
    row_bytes = src[0].nbytes if src.shape[0] else 0
    dst = np.zeros((idx.shape[0],) + src.shape[1:], dtype=src.dtype)
    rc = lib().oracle_gather_rows(_p(src), ctypes.c_int64(row_bytes), _p(idx),
                                  ctypes.c_int64(idx.shape[0]), ctypes.c_int64(src.shape[0]), _p(dst))
    if rc != 0:
        raise RuntimeError("oracle_gather_rows rc=%d" % rc)
    return dst


def vtrace(v_tm1, v_t, r_t, disc, rho, lambda_=1.0, clip_rho=1.0, clip_pg_rho=1.0):
    """Batch-major ``[B, T-1]`` float32 contiguous arrays -> (adv, err, q)."""
    arrs = [np.ascontiguousarray(x, dtype=np.float32) for x in (v_tm1, v_t, r_t, disc, rho)]
    B, Tm = arrs[0].shape
    outs = [np.zeros((B, Tm), dtype=np.float32) for _ in range(3)]
    lib().oracle_vtrace(*[_p(x) for x in arrs], ctypes.c_int64(B), ctypes.c_int64(Tm),
                        ctypes.c_int64(Tm), ctypes.c_int64(1), ctypes.c_float(lambda_),
                        ctypes.c_float(clip_rho), ctypes.c_float(clip_pg_rho),
                        *[_p(x) for x in outs])
    return tuple(outs)


def lambda_returns(r_t, disc, v_t, lambda_):
    arrs = [np.ascontiguousarray(x, dtype=np.float32) for x in (r_t, disc, v_t)]
    B, T = arrs[0].shape
    out = np.zeros((B, T), dtype=np.float32)
    lib().oracle_lambda_returns(*[_p(x) for x in arrs], ctypes.c_int64(B), ctypes.c_int64(T),
                                ctypes.c_int64(T), ctypes.c_int64(1), ctypes.c_float(lambda_), _p(out))
    return out


def gae(r_t, disc, values, lambda_):
    """``r_t, disc [B,T]``, ``values [B,T+1]`` -> ``[B,T]``."""
    arrs = [np.ascontiguousarray(x, dtype=np.float32) for x in (r_t, disc, values)]
    B, T = arrs[0].shape
    out = np.zeros((B, T), dtype=np.float32)
    lib().oracle_gae(*[_p(x) for x in arrs], ctypes.c_int64(B), ctypes.c_int64(T),
                     ctypes.c_int64(T), ctypes.c_int64(1), ctypes.c_int64(T + 1), ctypes.c_int64(1),
                     ctypes.c_float(lambda_), _p(out))
    return out
