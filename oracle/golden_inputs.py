"""Seeded inputs shared by oracle/gen_golden.py and the tests (TEST INFRASTRUCTURE ONLY).

The large fused-GEMV cases commit only the reference's OUTPUT plus a sha256 of the inputs; both
sides regenerate the inputs from the seed with this function (no TVM / reference import here).
"""
import hashlib

import numpy as np

from . import quant_oracle as qo


def sha(*arrs):
    h = hashlib.sha256()
    for a in arrs:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def synth_fused_inputs(fmt, K, N, seed):
    rng = np.random.default_rng(seed)
    R, G = fmt.n_words(K), fmt.n_groups(K)
    packed = rng.integers(0, 2**32, size=(R, N), dtype=np.uint64).astype(np.uint32)
    if fmt.nbit == 3:
        packed &= np.uint32(0x3FFFFFFF)
    scale = (rng.random((G, N), dtype=np.float32) * 0.01).astype(fmt.np_dtype)
    x = rng.standard_normal(K, dtype=np.float32).astype(fmt.np_dtype)
    ep = rng.standard_normal(N, dtype=np.float32).astype(fmt.np_dtype)
    return packed, scale, x, ep


def epilogue_of(prim_func_name: str) -> int:
    """Epilogue encoded in a reference PrimFunc name (decode_matmul_ewise.py:78-106 naming)."""
    for key, e in [("_add", "add"), ("_multiply", "multiply"), ("_silu", "silu"), ("_cast", "cast_f32")]:
        if key in prim_func_name:
            return qo.EPILOGUES[e]
    return qo.EP_NONE
