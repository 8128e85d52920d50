"""Operator-by-operator check of a decode-step launch sequence against the oracle (TEST INFRASTRUCTURE ONLY).

Used by tests/, __graft_entry__.smoke() and bench.py's `parity` record (outside every timed region): the product
(`mlc_llm_b200.decode_stack.DecodeStack.snapshot()`) hands over, for every fused dequant-GEMV it launched, the
host copies of the activation it READ and the output it WROTE; each operator is then re-evaluated here in f64 from
the same packed weights and the same input bits and compared under the stated per-operator tolerance.  Checking
operator by operator (instead of only the final logits) keeps the bound rigorous: no error is allowed to hide
behind the amplification of a 32-layer chain.

Reference semantics of one operator: mlc_llm/dispatch/dispatch_tir_operator_adreno.py:5-23 (matmul) and :86-91,
:150-152, :216-228, :364-369 (epilogues); gpt_neox/dolly_v2_3b_mod.py:41-46, 204-251 (bias, bias + GELU).
"""
from __future__ import annotations

import math

import numpy as np

from . import c_oracle as co
from . import quant_oracle as qo


def _erf(v):
    return np.vectorize(math.erf, otypes=[np.float64])(v)


def op_truth_and_tol(packed, scale, x, fmt: qo.Fmt, epilogue: int, ep0, K: int, group_scaled: bool):
    """(f64 truth of the fused operator, elementwise tolerance).

    The matmul tolerance t is SURVEY.md 8c(ii) (oracle.quant_oracle.tolerance) - for the tcgen05 decode program
    (`group_scaled`) the bound against the UNROUNDED weights with its cancellation term
    (quant_oracle.tolerance_group_scaled).  Epilogues are evaluated by the kernels op for op in the model dtype
    (f16: every op rounds, 2^-11 relative), so the bound is t pushed through the epilogue's derivative plus one
    rounding per f16 operation of the epilogue."""
    f32 = fmt.dtype == "float32"
    out_f32 = f32 or epilogue == qo.EP_CAST_F32
    if group_scaled:
        y64, ab, cc = co.gemv_f64_unrounded(packed, scale, x, fmt, K)
        t = qo.tolerance_group_scaled(y64, ab, cc, out_f32)
    else:
        y64, ab = co.gemv_f64(packed, scale, x, fmt, K)
        t = qo.tolerance(y64, ab, out_f32)
    u = 2.0**-24 if f32 else 2.0**-11  # one rounding in the model dtype
    e = None if ep0 is None else ep0.astype(np.float64).reshape(-1)
    if epilogue in (qo.EP_NONE, qo.EP_CAST_F32):
        return y64, t
    if epilogue in (qo.EP_ADD, qo.EP_BIAS):
        out = e + y64
        return out, 1.001 * t + u * np.abs(out) + 2.0**-24
    if epilogue == qo.EP_MULTIPLY:
        out = e * y64
        return out, 1.001 * np.abs(e) * t + u * np.abs(out) + 2.0**-24
    if epilogue == qo.EP_SILU:
        out = y64 / (1.0 + np.exp(-y64))
        # y * (1 / (1 + exp(-y))): four dtype roundings; |d silu / dy| <= 1.1; exp(-y) overflows f16 for y < -11.09
        # (the f16 chain then returns 0 where the true value is |y| e^y <= 2e-4)
        tol = 1.1 * t + 4 * u * np.abs(out) + 2.0**-24
        if not f32:
            tol = tol + np.where(y64 < -10.5, 2.5e-4, 0.0)
        return out, tol
    if epilogue == qo.EP_BIAS_GELU:
        v = e + y64
        out = v * (0.5 + 0.5 * _erf(v * 0.70710678118654757))
        # v = y + b; v * (0.5 + erf(v / sqrt 2) * 0.5): |d gelu / dv| <= 1.13; the factor h = 0.5 + 0.5 erf is formed in
        # dtype (absolute error ~2 u), so the product carries |v| * 2 u on top of the relative roundings
        tol = 1.13 * (t + u * np.abs(v)) + 2 * u * np.abs(v) + 4 * u * np.abs(out) + 2.0**-24
        return out, tol
    raise ValueError(f"epilogue {epilogue}")


def check_launches(launches, weights, fmt: qo.Fmt, group_scaled: bool):
    """launches: DecodeStack.snapshot() output (host numpy arrays, full / unsharded tensors);
    weights(layer, name) -> (packed uint32[R, N], scale[G, N]) host arrays of the FULL (unsharded) Linear.
    Returns the `parity` record."""
    worst, worst_op, n_ops, n_vals = 0.0, None, 0, 0
    for ln in launches:
        x = ln["x"]
        if ln["x_mul"] is not None:
            x = (x * ln["x_mul"]).astype(x.dtype)  # the kernels form x * x_mul with one rounding in dtype
        for op in ln["ops"]:
            packed, scale = weights(op["layer"], op["name"])
            K = x.size
            truth, tol = op_truth_and_tol(packed, scale, x, fmt, qo.EPILOGUES[op["epilogue"]], op["ep0"], K, group_scaled)
            got = op["out"].astype(np.float64).reshape(-1)
            if not np.isfinite(got).all():
                return {"ok": False, "max_err_over_tol": float("inf"), "worst_op": f"L{op['layer']}.{op['name']} (non-finite)",
                        "ops_checked": n_ops}
            r = float((np.abs(got - truth) / tol).max())
            n_ops += 1
            n_vals += got.size
            if r > worst:
                worst, worst_op = r, f"L{op['layer']}.{op['name']}+{op['epilogue']}"
    return {"ok": bool(worst <= 1.0), "max_err_over_tol": round(worst, 4), "worst_op": worst_op, "ops_checked": n_ops,
            "values_checked": n_vals,
            "bound": ("|y - y64u| <= rel|y| + 2^-14 sum|xw| + 2^-22 sum off|x|s + 2^-24 vs f64 on the unrounded (q-7)s weights"
                      if group_scaled else "|y - y64| <= rel|y| + 2^-14 sum|xw| + 2^-24 vs f64 on the reference's dequantized weights")
                     + "; epilogues: the bound pushed through the op + one dtype rounding per op"}
