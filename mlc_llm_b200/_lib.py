"""ctypes binding of libb200q.so (the C-ABI declared in include/b200q.h).

The library is built in-tree by ``make -C mlc_llm_b200/csrc`` (or ``__graft_entry__.build()``).
There is no fallback of any kind: if the shared object is missing or a call fails, an exception
is raised.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200Q_LIB: another build of the SAME library (kernel experiments: EXTRA=-D... OUT=...); never a fallback
LIB_PATH = os.environ.get("B200Q_LIB") or os.path.join(_HERE, "libb200q.so")
CSRC = os.path.join(_HERE, "csrc")

c_void_p, c_int, c_uint, c_ll, c_ull = (
    ctypes.c_void_p,
    ctypes.c_int,
    ctypes.c_uint,
    ctypes.c_longlong,
    ctypes.c_ulonglong,
)


class GemmOp(ctypes.Structure):
    """b200q_gemm_op (include/b200q.h)."""

    _fields_ = [
        ("packed", c_void_p),
        ("scale", c_void_p),
        ("zmin", c_void_p),
        ("ep0", c_void_p),
        ("ep1", c_void_p),
        ("ep2", c_void_p),
        ("Y", c_void_p),
        ("N", c_int),
        ("epilogue", c_int),
    ]


class GemvOp(ctypes.Structure):
    """b200q_gemv_op (include/b200q.h)."""

    _fields_ = [
        ("packed", c_void_p),
        ("scale", c_void_p),
        ("ep0", c_void_p),
        ("out", c_void_p),
        ("N", c_int),
        ("epilogue", c_int),
    ]


class GemvStep(ctypes.Structure):
    """b200q_gemv_step (include/b200q.h)."""

    _fields_ = [("ops", GemvOp * 4), ("n_ops", c_int), ("x", c_void_p), ("x_mul", c_void_p), ("K", c_int)]


# name -> (restype, argtypes); must list every symbol include/b200q.h declares
PROTOTYPES = {
    "b200q_version": (c_int, []),
    "b200q_last_error": (ctypes.c_char_p, []),
    "b200q_launch_count": (c_ull, []),
    "b200q_n_groups": (c_int, [c_int, c_int]),
    "b200q_n_words": (c_int, [c_int, c_int]),
    "b200q_gemv_bytes": (c_ll, [c_int, c_int, c_int, c_int]),
    "b200q_gemv_workspace_bytes": (ctypes.c_size_t, []),
    "b200q_gemv": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_uint, c_void_p, ctypes.c_size_t, c_void_p]),
    "b200q_gemv_static": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_uint, c_void_p]),
    "b200q_gemv_multi": (c_int, [ctypes.POINTER(GemvOp), c_int, c_void_p, c_void_p, c_int, c_int, c_uint, c_void_p,
                                 ctypes.c_size_t, c_void_p]),
    "b200q_gemv_program_bytes": (ctypes.c_size_t, [c_int]),
    "b200q_gemv_program_build": (c_int, [ctypes.POINTER(GemvStep), c_int, c_int, c_void_p, ctypes.c_size_t]),
    "b200q_gemv_program_workspace_bytes": (ctypes.c_size_t, [c_void_p]),
    "b200q_gemv_program_run": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.c_size_t, c_void_p]),
    "b200q_gemv_program_build_tp": (c_int, [ctypes.POINTER(GemvStep), c_int, c_int, c_void_p, c_int, c_int, c_void_p,
                                            ctypes.c_size_t]),
    "b200q_gemv_program_run_tp": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.c_size_t, c_void_p, c_void_p]),
    "b200q_gemv_program_set_watchdog": (None, [c_void_p, c_ll]),
    "b200q_gemv_tune": (None, [c_int, c_int, c_int]),
    "b200q_gemv_host": (c_int, [c_void_p] * 7 + [c_int] * 4 + [c_void_p, ctypes.c_size_t, c_void_p]),
    "b200q_dequant": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_void_p]),
    "b200q_take_decode": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_void_p]),
    "b200q_encode": (c_int, [c_void_p] * 3 + [c_int] * 4 + [c_void_p]),
    "b200q_gemv_asym": (c_int, [c_void_p] * 6 + [c_int] * 4 + [c_uint, c_void_p, ctypes.c_size_t, c_void_p]),
    "b200q_dequant_asym": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_void_p]),
    "b200q_take_decode_asym": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p]),
    "b200q_encode_asym": (c_int, [c_void_p] * 4 + [c_int] * 4 + [c_void_p]),
    "b200q_gemm_asym": (c_int, [c_void_p] * 8 + [c_int] * 5 + [c_uint, c_void_p]),
    "b200q_gemm": (c_int, [c_void_p] * 7 + [c_int] * 5 + [c_uint, c_void_p]),
    "b200q_gemm_workspace_bytes": (ctypes.c_size_t, []),
    "b200q_debug_gemm_schedule": (c_int, [c_int] * 6 + [c_void_p, c_int]),
    "b200q_debug_gemv_schedule": (c_int, [c_int] * 5 + [c_void_p, c_int, c_void_p]),
    "b200q_gemm_multi": (c_int, [ctypes.POINTER(GemmOp), c_int, c_void_p, c_int, c_int, c_int, c_uint, c_void_p, ctypes.c_size_t,
                                 c_void_p]),
    "b200q_gemm_ws": (c_int, [c_void_p] * 8 + [c_int] * 5 + [c_uint, c_void_p, ctypes.c_size_t, c_void_p]),
    "b200q_nt_matmul": (c_int, [c_void_p] * 6 + [c_int] * 4 + [c_void_p]),
    "b200q_nt_matmul_ws": (c_int, [c_void_p] * 6 + [c_int] * 4 + [c_uint, c_void_p, ctypes.c_size_t, c_void_p]),
}

_lib = None


class B200QError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"b200q error {code}: {msg}")
        self.code = code


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a into mlc_llm_b200/libb200q.so (nvcc cross-compiles
    without a GPU)."""
    cmd = ["make", "-C", CSRC, "-j8"] + (["-B"] if force else [])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise RuntimeError("building libb200q.so failed")
    return LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C mlc_llm_b200/csrc` "
                "(there is no CPU / PyTorch fallback for this path)"
            )
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the header and the library drift apart
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise B200QError(code, lib().b200q_last_error().decode())
