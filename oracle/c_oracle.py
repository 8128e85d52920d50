"""ctypes binding of oracle/_build/liboracle.so  (TEST INFRASTRUCTURE ONLY, see oracle/oracle.c)."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

from . import quant_oracle as qo

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = ctypes.CDLL(_SO)
        _lib.orc_num_threads.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _suf(fmt: qo.Fmt) -> str:
    return "f16" if fmt.dtype == "float16" else "f32"


def num_threads() -> int:
    return lib().orc_num_threads()


def set_threads(n: int):
    lib().orc_set_threads(int(n))


def decode(packed, scale, K, fmt: qo.Fmt, transpose_output=False):
    N = packed.shape[1]
    out = np.empty((N, K) if transpose_output else (K, N), fmt.np_dtype)
    getattr(lib(), "orc_decode_" + _suf(fmt))(
        _p(packed), _p(scale), _p(out), K, N, fmt.nbit, fmt.group, int(transpose_output)
    )
    return out


def gemv_ref(packed, scale, x, fmt: qo.Fmt, epilogue=qo.EP_NONE, ep0=None, K=None, reassociate=False):
    """The reference's dtype-accumulating GEMV (bit-identical to its TVM C-backend build)."""
    N = packed.shape[1]
    K = x.size if K is None else K
    packed = np.ascontiguousarray(packed)
    scale = np.ascontiguousarray(scale)
    x = np.ascontiguousarray(x.reshape(-1))
    if ep0 is not None:
        ep0 = np.ascontiguousarray(ep0.reshape(-1).astype(fmt.np_dtype))
    odt = np.float32 if epilogue == qo.EP_CAST_F32 else fmt.np_dtype
    out = np.empty(N, odt)
    getattr(lib(), "orc_gemv_ref_" + _suf(fmt))(
        _p(packed), _p(scale), _p(x), _p(ep0), _p(out), K, N, fmt.nbit, fmt.group, int(epilogue), int(reassociate)
    )
    return out


def gemv_f64(packed, scale, x, fmt: qo.Fmt, K=None):
    N = packed.shape[1]
    K = x.size if K is None else K
    y = np.empty(N, np.float64)
    a = np.empty(N, np.float64)
    x = np.ascontiguousarray(x.reshape(-1))
    getattr(lib(), "orc_gemv_f64_" + _suf(fmt))(
        _p(np.ascontiguousarray(packed)), _p(np.ascontiguousarray(scale)), _p(x), _p(y), _p(a), K, N, fmt.nbit, fmt.group
    )
    return y, a


def gemv_f64_unrounded(packed, scale, x, fmt: qo.Fmt, K=None):
    """C version of quant_oracle.gemv_f64_unrounded (f16 formats): (y, sum|x w|, cancellation magnitude)."""
    assert fmt.dtype == "float16"
    N = packed.shape[1]
    K = x.size if K is None else K
    y, a, c = np.empty(N, np.float64), np.empty(N, np.float64), np.empty(N, np.float64)
    x = np.ascontiguousarray(x.reshape(-1))
    lib().orc_gemv_f64u_f16(_p(np.ascontiguousarray(packed)), _p(np.ascontiguousarray(scale)), _p(x), _p(y), _p(a), _p(c),
                            K, N, fmt.nbit, fmt.group)
    return y, a, c


def decode_asym(packed, scale, mn, K, fmt: qo.Fmt, transpose_output=False):
    """C version of quant_oracle.decode_asym (mn is None for the f32 scheme: scale holds bf16 scale | bf16 min)."""
    N = packed.shape[1]
    out = np.empty((N, K) if transpose_output else (K, N), fmt.np_dtype)
    getattr(lib(), "orc_decode_asym_" + _suf(fmt))(_p(np.ascontiguousarray(packed)), _p(np.ascontiguousarray(scale)),
                                                  _p(None if mn is None else np.ascontiguousarray(mn)), _p(out), K, N,
                                                  int(transpose_output))
    return out


def gemv_ref_asym(packed, scale, mn, x, fmt: qo.Fmt, epilogue=qo.EP_NONE, ep0=None):
    N = packed.shape[1]
    x = np.ascontiguousarray(x.reshape(-1))
    if ep0 is not None:
        ep0 = np.ascontiguousarray(ep0.reshape(-1).astype(fmt.np_dtype))
    out = np.empty(N, np.float32 if epilogue == qo.EP_CAST_F32 else fmt.np_dtype)
    getattr(lib(), "orc_gemv_ref_asym_" + _suf(fmt))(_p(np.ascontiguousarray(packed)), _p(np.ascontiguousarray(scale)),
                                                    _p(None if mn is None else np.ascontiguousarray(mn)), _p(x), _p(ep0),
                                                    _p(out), x.size, N, int(epilogue))
    return out


def gemv_f64_asym(packed, scale, mn, x, fmt: qo.Fmt):
    N = packed.shape[1]
    x = np.ascontiguousarray(x.reshape(-1))
    y, a = np.empty(N, np.float64), np.empty(N, np.float64)
    getattr(lib(), "orc_gemv_f64_asym_" + _suf(fmt))(_p(np.ascontiguousarray(packed)), _p(np.ascontiguousarray(scale)),
                                                    _p(None if mn is None else np.ascontiguousarray(mn)), _p(x), _p(y), _p(a),
                                                    x.size, N)
    return y, a


def nt_matmul_ref_f16(X, T):
    M, K = X.shape
    N = T.shape[0]
    Y = np.empty((M, N), np.float16)
    lib().orc_nt_matmul_ref_f16(_p(np.ascontiguousarray(X)), _p(np.ascontiguousarray(T)), _p(Y), M, K, N)
    return Y


def nt_matmul_f64_f16(X, T):
    M, K = X.shape
    N = T.shape[0]
    Y = np.empty((M, N), np.float64)
    A = np.empty((M, N), np.float64)
    lib().orc_nt_matmul_f64_f16(_p(np.ascontiguousarray(X)), _p(np.ascontiguousarray(T)), _p(Y), _p(A), M, K, N)
    return Y, A
