"""numpy/ctypes front-end of oracle.c (GAE + interval algorithms on the CPU). TEST INFRASTRUCTURE.

Pinned by tests/golden/gae_*.npz and interval_*.npz, which gen_golden.py produced by running the
real reference (pygae1d_nolp_misalign, _slice_intervals_py, _set_intervals_py, merge_intervals).
"""

from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")


def build(force: bool = False) -> str:
    """Compile oracle.c (and oracle/_ref when /root/reference exists) with oracle/Makefile."""
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(
        os.path.join(_HERE, "oracle.c")
    ):
        subprocess.run(["make", "-C", _HERE, "liboracle.so"], check=True, capture_output=True)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_LIB)
        _lib.oracle_slice_intervals.restype = ctypes.c_int64
        _lib.oracle_set_intervals.restype = ctypes.c_int64
        _lib.oracle_merge_intervals.restype = ctypes.c_int64
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(ctypes.c_void_p)


def gae_py(rewards, values, cu_seqlens, bootstrap, gamma: float, lam: float):
    """pygae1d_nolp_misalign semantics (ppo_functional.py:292-323). Inputs are not modified."""
    r = np.ascontiguousarray(rewards, dtype=np.float32)
    v = np.ascontiguousarray(values, dtype=np.float32)
    cu = np.ascontiguousarray(cu_seqlens, dtype=np.int32)
    b = np.ascontiguousarray(bootstrap, dtype=np.float32)
    assert v.shape[0] == r.shape[0] + cu.shape[0] - 1
    adv, ret = np.empty_like(r), np.empty_like(r)
    lib().oracle_gae_py(_p(r), _p(v), _p(cu), _p(b), ctypes.c_int(cu.shape[0] - 1),
                        ctypes.c_double(gamma), ctypes.c_double(lam), _p(adv), _p(ret))
    return adv, ret


def gae_cuda_order(rewards, values, cu_seqlens, bootstrap, gamma: float, lam: float):
    """csrc/cugae/gae.cu:10-28 semantics (fp32 FMA contraction, select on bootstrap)."""
    r = np.ascontiguousarray(rewards, dtype=np.float32)
    v = np.ascontiguousarray(values, dtype=np.float32)
    cu = np.ascontiguousarray(cu_seqlens, dtype=np.int32)
    b = np.ascontiguousarray(np.asarray(bootstrap) != 0, dtype=np.uint8)
    adv, ret = np.empty_like(r), np.empty_like(r)
    lib().oracle_gae_cuda(_p(r), _p(v), _p(cu), _p(b), ctypes.c_int(cu.shape[0] - 1),
                          ctypes.c_float(gamma), ctypes.c_float(lam), _p(adv), _p(ret))
    return adv, ret


def gae_2d_olp(rewards, values, done, done_y_indices, cu_num_dones, is_truncate, gamma: float, lam: float):
    """csrc/cugae/gae.cu:62-92 (gae_kernel_2d_olp) on [B,H] / [B,H+1] arrays."""
    r = np.ascontiguousarray(rewards, dtype=np.float32)
    v = np.ascontiguousarray(values, dtype=np.float32)
    d = np.ascontiguousarray(np.asarray(done) != 0, dtype=np.uint8)
    y = np.ascontiguousarray(done_y_indices, dtype=np.int32)
    cu = np.ascontiguousarray(cu_num_dones, dtype=np.int32)
    tr = np.ascontiguousarray(np.asarray(is_truncate) != 0, dtype=np.uint8)
    B, H = r.shape
    assert v.shape == (B, H + 1) and d.shape == (B, H + 1) and cu.shape == (B + 1,)
    adv, ret = np.zeros_like(r), np.zeros_like(r)
    lib().oracle_gae_2d_olp(_p(r), _p(v), _p(d), _p(y), _p(cu), _p(tr), ctypes.c_int(B), ctypes.c_int(H),
                            ctypes.c_float(gamma), ctypes.c_float(lam), _p(adv), _p(ret))
    return adv, ret


def gae_f64(rewards, values, cu_seqlens, bootstrap, gamma: float, lam: float):
    r = np.ascontiguousarray(rewards, dtype=np.float32)
    v = np.ascontiguousarray(values, dtype=np.float32)
    cu = np.ascontiguousarray(cu_seqlens, dtype=np.int32)
    b = np.ascontiguousarray(np.asarray(bootstrap) != 0, dtype=np.uint8)
    adv, ret = np.empty(r.shape, np.float64), np.empty(r.shape, np.float64)
    lib().oracle_gae_f64(_p(r), _p(v), _p(cu), _p(b), ctypes.c_int(cu.shape[0] - 1),
                         ctypes.c_double(gamma), ctypes.c_double(lam), _p(adv), _p(ret))
    return adv, ret


def gae_py_loops(rewards, values, cu_seqlens, bootstrap, gamma: float, lam: float):
    """The same recurrence as python loops over numpy float32 scalars (small cases only);
    an independent check of oracle.c's rounding sequence."""
    r = np.asarray(rewards, np.float32)
    v = np.asarray(values, np.float32).copy()
    cu = np.asarray(cu_seqlens, np.int64)
    f = np.float32
    adv = np.zeros_like(r)
    ret = np.zeros_like(r)
    c = f(gamma * lam)
    for i in range(len(cu) - 1):
        rs, re = int(cu[i]), int(cu[i + 1])
        vs = rs + i
        last = f(0)
        for t in range(re - rs - 1, -1, -1):
            nv = v[vs + t + 1]
            if t == re - rs - 1:
                nv = f(nv * f(bootstrap[i]))
            d = f(f(r[rs + t] + f(f(gamma) * nv)) - v[vs + t])
            last = f(d + f(c * last))
            adv[rs + t] = last
            ret[rs + t] = f(last + v[vs + t])
    return adv, ret


def slice_intervals(src: np.ndarray, intervals) -> np.ndarray:
    """_slice_intervals_py (flatten_param.py:65-74) on raw bytes: bit-exact for any dtype."""
    src = np.ascontiguousarray(src)
    iv = np.ascontiguousarray(np.asarray(intervals, dtype=np.int64).reshape(-1, 2))
    total = int((iv[:, 1] - iv[:, 0]).sum()) if len(iv) else 0
    out = np.empty(total, dtype=src.dtype)
    n = lib().oracle_slice_intervals(_p(src), ctypes.c_int64(src.shape[0]), _p(out), _p(iv),
                                     ctypes.c_int64(len(iv)), ctypes.c_int(src.itemsize))
    assert n == total, "interval out of range"
    return out


def set_intervals(src: np.ndarray, dst: np.ndarray, intervals) -> None:
    """_set_intervals_py (flatten_param.py:77-90); writes into ``dst`` in place."""
    src = np.ascontiguousarray(src)
    assert dst.flags.c_contiguous and dst.dtype == src.dtype
    iv = np.ascontiguousarray(np.asarray(intervals, dtype=np.int64).reshape(-1, 2))
    n = lib().oracle_set_intervals(_p(src), _p(dst), ctypes.c_int64(dst.shape[0]), _p(iv),
                                   ctypes.c_int64(len(iv)), ctypes.c_int(src.itemsize))
    assert n == src.shape[0], (n, src.shape[0])  # flatten_param.py:90


def merge_intervals(intervals):
    iv = np.ascontiguousarray(np.asarray(intervals, dtype=np.uint64).reshape(-1, 2))
    out = np.empty_like(iv)
    m = lib().oracle_merge_intervals(_p(iv), ctypes.c_int64(len(iv)), _p(out))
    return [(int(a), int(b)) for a, b in out[:m]]
