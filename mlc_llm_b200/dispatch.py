"""Dispatch registry: the plug-in seam of the reference, restated for the CUDA path.

The reference replaces a compiler-generated PrimFunc by a hand-scheduled one when the two are
structurally equal (mlc_llm/dispatch/dispatch_tir_operator.py:22-33; lookup_func,
mlc_llm/dispatch/llama/main.py:6452-6458).  A key there is one exact
(quantization scheme, K, N, epilogue) tuple; a miss returns None and the caller keeps the
unscheduled function.  ``lookup`` keeps that contract: it returns a destination-passing callable
``f(packed, scale, x[, ep0...], out)`` - the PrimFunc argument order
(dispatch_tir_operator_adreno.py:6,128) - or None when this library does not handle the tuple.
There is no fallback behind a miss: callers decide, nothing silently runs on the CPU.
"""
from __future__ import annotations

from typing import Callable, Optional

from . import ops
from .quant_spec import EPILOGUES, SPECS, QuantSpec

# what the reference's FuseDecodeMatmulEwise / FuseOps produce for each path
GEMV_EPILOGUES = ("none", "add", "multiply", "silu", "cast_f32", "bias", "bias_gelu")
GEMM_EPILOGUES = ("none", "add", "multiply", "silu", "cast_f32", "bias", "bias_gelu", "bias_add2")
N_OPERANDS = {"none": 0, "add": 1, "multiply": 1, "silu": 0, "cast_f32": 0, "bias": 1, "bias_gelu": 1, "bias_add2": 3}


def reference_name(spec: QuantSpec, K: int, N: int, epilogue: str, m_is_one: bool) -> str:
    """A readable stand-in for the reference's generated function names
    (fused_decode4_matmul3, fused_decode5_fused_matmul6_silu1, fused_decode_NT_matmul ...)."""
    core = "matmul" if m_is_one else "NT_matmul"
    ep = "" if epilogue == "none" else "_" + epilogue
    return f"fused_decode_{spec.name}_{core}_k{K}_n{N}{ep}"


def lookup(spec, K: int, N: int, epilogue: str = "none", m_is_one: bool = True) -> Optional[Callable]:
    """Return the CUDA operator for (scheme, K, N, epilogue) or None (the reference's miss)."""
    if isinstance(spec, str):
        spec = SPECS.get(spec)
    if spec is None or epilogue not in EPILOGUES:
        return None
    if K <= 0 or N <= 0 or N % 8 != 0:
        return None
    if m_is_one:
        if epilogue not in GEMV_EPILOGUES:
            return None

        def gemv_op(packed, scale, *rest):
            zmin = None
            if spec.has_min:  # (packed, scale, min, x[, operand], out): dispatch/llama/main.py:4979
                zmin, *rest = rest
            x, *rest = rest
            *eps, out = rest
            if len(eps) != N_OPERANDS[epilogue]:
                raise TypeError(f"{reference_name(spec, K, N, epilogue, True)} takes {N_OPERANDS[epilogue]} epilogue operand(s)")
            return ops.gemv(packed, scale, x, spec, epilogue, eps[0] if eps else None, out=out, pdl=False, zmin=zmin)

        gemv_op.__name__ = reference_name(spec, K, N, epilogue, True)
        return gemv_op
    # prefill: dequantize inside the tensor-core GEMM (f16 schemes with uint32 integer storage only; the others take
    # the reference's own route, lookup_decode + a dense NT matmul)
    if spec.dtype != "float16" or spec.is_variant or epilogue not in GEMM_EPILOGUES:
        return None

    def gemm_op(packed, scale, *rest):
        zmin = None
        if spec.has_min:
            zmin, *rest = rest
        x, *rest = rest
        *eps, out = rest
        if len(eps) != N_OPERANDS[epilogue]:
            raise TypeError(f"{reference_name(spec, K, N, epilogue, False)} takes {N_OPERANDS[epilogue]} epilogue operand(s)")
        eps = list(eps) + [None] * (3 - len(eps))
        return ops.gemm(packed, scale, x, spec, epilogue, eps[0], eps[1], eps[2], out=out, zmin=zmin)

    gemm_op.__name__ = reference_name(spec, K, N, epilogue, False)
    return gemm_op


def lookup_decode(spec, K: int, N: int, transpose_output: bool) -> Optional[Callable]:
    """decode* (dequantize-to-dense) operators: f(packed, scale, out)."""
    if isinstance(spec, str):
        spec = SPECS.get(spec)
    if spec is None or N % 8 != 0:
        return None

    def decode_op(packed, scale, *rest):
        zmin = None
        if spec.has_min:  # decode4_fp16(data, scale, min, out): dispatch/llama/main.py:4042
            zmin, *rest = rest
        (out,) = rest
        return ops.dequantize(packed, scale, K, spec, transpose_output, out=out, zmin=zmin)

    decode_op.__name__ = f"decode_{spec.name}_k{K}_n{N}" + ("_T" if transpose_output else "")
    return decode_op


def lookup_program(spec, call_sites) -> Optional["ops.GemvProgram"]:
    """The whole fused-GEMV chain of one `decoding` call as ONE operator.

    The reference dispatches PrimFunc by PrimFunc (lookup_func, mlc_llm/dispatch/llama/main.py:6452-6458); a
    B200-native drop-in can additionally take the LIST of consecutive fused_decode*_matmul* call sites of the
    compiled `decoding` function (mlc_llm/relax_model/llama.py:581-611) and run them as one persistent launch.
    call_sites: [{"weights": [(packed, scale), ...], "x": tensor, "x_mul": tensor | None,
                  "epilogues": [...], "ep0s": [...], "outs": [...]}, ...]   (ops.GemvProgram's step format:
    every entry = the 1..4 GEMVs that read one activation).  Consecutive call sites only: a non-GEMV kernel
    between two of them ends the list (DecodeStack(segments="reference") shows the cut points).
    Returns None - the reference's miss - when any call site is outside what `lookup` handles."""
    if isinstance(spec, str):
        spec = SPECS.get(spec)
    if spec is None or not call_sites:
        return None
    for cs in call_sites:
        K = cs["x"].shape[-1]
        eps = cs.get("epilogues") or ["none"] * len(cs["weights"])
        if not (1 <= len(cs["weights"]) <= 4):
            return None
        for (packed, _), ep in zip(cs["weights"], eps):
            if lookup(spec, K, packed.shape[1], ep, m_is_one=True) is None:
                return None
    return ops.GemvProgram(list(call_sites), spec)
