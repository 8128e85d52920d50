"""CPU oracle for the group-quantized dequantize+matmul path  (TEST INFRASTRUCTURE ONLY).

This file is the checker, never the product: only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import it.  The shipped
path (``mlc_llm_b200``) never imports anything from ``oracle/``.

It is a numpy restatement of the reference's TVM tensor-expression formulas.  Every
function cites the reference lines (relative to /root/reference) it follows.

Parity pin: ``oracle/gen_golden.py`` executes the reference's *own* ``encoding_func`` /
``decoding_func`` / ``fused_decode*_matmul*`` PrimFuncs (TVM C backend, this container) and
stores their outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks this
restatement bit-for-bit against those vectors.  The reference itself ships no golden vectors
(SURVEY.md section 4), so those fixtures are the pin.
"""
from __future__ import annotations

import dataclasses
import math

import numpy as np

# --------------------------------------------------------------------------------------
# format description  (reference flags: build.py:19-25,114; utils.py:31-36)
# --------------------------------------------------------------------------------------


@dataclasses.dataclass(frozen=True)
class Fmt:
    """One group-quantization scheme ("_0" = symmetric; "_asym" = the asymmetric scheme of te_encode_asym).

    name          BASELINE name
    nbit          bits per value                       (quantization.py:313  nbit = int(mode[-1]))
    group         values sharing one scale             (build.py:114  40 if int3 else 32)
    storage_nbit  bits of the storage word             (build.py:22-25)
    dtype         model dtype = scale / activation / accumulate dtype (utils.py:31-36)
    """

    name: str
    nbit: int
    group: int
    dtype: str
    storage_nbit: int = 32
    sym: bool = True  # --quantization-sym; False = the reference CLI's default (asymmetric: scale + min per group)

    @property
    def npu(self) -> int:  # values per storage word  (quantization.py:144)
        return self.storage_nbit // self.nbit

    @property
    def maxint(self) -> int:  # quantization.py:145
        return (1 << (self.nbit - 1)) - 1

    @property
    def np_dtype(self):
        return np.float16 if self.dtype == "float16" else np.float32

    def n_groups(self, K: int) -> int:  # quantization.py:143
        return -(-K // self.group)

    def n_words(self, K: int) -> int:  # quantization.py:170  n_i32
        return -(-self.group // self.npu) * self.n_groups(K)


FORMATS = {
    "q4f16_0": Fmt("q4f16_0", 4, 32, "float16"),
    "q3f16_0": Fmt("q3f16_0", 3, 40, "float16"),
    "q4f32_0": Fmt("q4f32_0", 4, 32, "float32"),
    "q8f16_0": Fmt("q8f16_0", 8, 32, "float16"),
}
# asymmetric int4 (build.py:19: --quantization-sym defaults to False): f16 = separate scale / min tensors,
# f32 = scale and min rounded to bf16 and packed into one uint32 (quantization.py:15-37,105-140,183-211)
ASYM_FORMATS = {
    "q4f16_asym": Fmt("q4f16_asym", 4, 32, "float16", sym=False),
    "q4f32_asym": Fmt("q4f32_asym", 4, 32, "float32", sym=False),
}
# the remaining --quantization-* combinations: int3 in uint16 words (5 values per word, the iOS
# build: ios/README.md:39; kernels dispatch/llama/main.py:4151-4203, 5815-6363) and fp4, an exponent-only 4-bit float
# (quantization.py:51-101), symmetric, uint32 words
VARIANT_FORMATS = {
    "q3f16_u16": Fmt("q3f16_u16", 3, 40, "float16", storage_nbit=16),
    "fp4f16_0": Fmt("fp4f16_0", 4, 32, "float16"),
    "fp4f32_0": Fmt("fp4f32_0", 4, 32, "float32"),
}
ALL_FORMATS = {**FORMATS, **ASYM_FORMATS, **VARIANT_FORMATS}


def is_fp4(fmt: "Fmt") -> bool:
    return fmt.name.startswith("fp4")


def storage_dtype(fmt: "Fmt"):
    return np.uint16 if fmt.storage_nbit == 16 else np.uint32

# epilogue ids shared with include/b200q.h
EP_NONE, EP_ADD, EP_MULTIPLY, EP_SILU, EP_CAST_F32, EP_BIAS, EP_BIAS_GELU, EP_BIAS_ADD2 = range(8)
EPILOGUES = {
    "none": EP_NONE,
    "add": EP_ADD,
    "multiply": EP_MULTIPLY,
    "silu": EP_SILU,
    "cast_f32": EP_CAST_F32,
    "bias": EP_BIAS,
    "bias_gelu": EP_BIAS_GELU,
    "bias_add2": EP_BIAS_ADD2,
}


def _round_half_away(x: np.ndarray) -> np.ndarray:
    """tir.round -> C roundf: halves away from zero (harness lowering; LLVM llvm.round)."""
    x64 = x.astype(np.float64)
    return np.copysign(np.floor(np.abs(x64) + 0.5), x64)


# --------------------------------------------------------------------------------------
# a1: encode   (transform/quantization.py:142-177  te_encode_sym, transpose=True)
# --------------------------------------------------------------------------------------


def encode_sym(W: np.ndarray, fmt: Fmt):
    """W[N, K] (fmt dtype) -> packed uint32[n_words, N], scale dtype[n_groups, N].

    Every arithmetic op is rounded to ``fmt.dtype`` exactly as the TE expression is typed.
    """
    dt = fmt.np_dtype
    assert W.dtype == dt and W.ndim == 2
    N, K = W.shape
    g, npu, maxint = fmt.group, fmt.npu, fmt.maxint
    G, R = fmt.n_groups(K), fmt.n_words(K)
    assert g % npu == 0  # quantization.py:146

    # max_abs over in-range k; out-of-range contributes tir.min_value(dtype)   (:150)
    Wp = np.full((N, G * g), np.finfo(dt).min, dtype=dt)
    Wp[:, :K] = np.abs(W)
    max_abs = Wp.reshape(N, G, g).max(axis=2)
    # scale = max(max_abs, 1e-4) / maxint, all in dtype                          (:152-156)
    scale = (np.maximum(max_abs, dt(1e-4)) / dt(maxint)).astype(dt)

    # q = clamp(round(W / scale + maxint), 0, 2*maxint)                          (:159-164)
    s_full = np.repeat(scale, g, axis=1)[:, :K]
    t = (W / s_full).astype(dt)
    t = (t + dt(maxint)).astype(dt)
    q = _round_half_away(t)
    q = np.minimum(np.maximum(q, 0.0), float(2 * maxint)).astype(np.uint32)

    # pack npu values per word, zero beyond K; word j of row n holds k = j*npu + t (:169-176)
    qp = np.zeros((N, R * npu), dtype=np.uint32)
    qp[:, :K] = q
    qp = qp.reshape(N, R, npu)
    packed = np.zeros((N, R), dtype=np.uint32)
    for tpos in range(npu):
        packed |= qp[:, :, tpos] << np.uint32(tpos * fmt.nbit)
    return np.ascontiguousarray(packed.T), np.ascontiguousarray(scale.T)


# --------------------------------------------------------------------------------------
# a2: decode   (quantization.py:213-230 te_decode_sym; :40-48 _tir_packed_uint_to_uint_to_float)
# --------------------------------------------------------------------------------------


def decode_sym(packed: np.ndarray, scale: np.ndarray, K: int, fmt: Fmt, transpose_output: bool = False):
    """packed uint32[R, N], scale dtype[G, N] -> w dtype[K, N]  (or [N, K] if transpose_output)."""
    dt = fmt.np_dtype
    assert packed.dtype == np.uint32 and scale.dtype == dt
    k = np.arange(K)
    word = packed[k // fmt.npu, :]
    shift = ((k % fmt.npu) * fmt.nbit).astype(np.uint32)[:, None]
    q = (word >> shift) & np.uint32((1 << fmt.nbit) - 1)
    w = (q.astype(dt) - dt(fmt.maxint)).astype(dt)  # exact in f16 and f32
    w = (w * scale[k // fmt.group, :]).astype(dt)  # the single rounding
    return np.ascontiguousarray(w.T) if transpose_output else w


# --------------------------------------------------------------------------------------
# quantized embedding gather  (quantization.py:235-270 te_take_decode_sym; transpose=False layout)
# --------------------------------------------------------------------------------------


def encode_sym_rowmajor(W: np.ndarray, fmt: Fmt):
    """Embedding layout: packed uint32[V, n_words], scale dtype[V, n_groups] (quantization.py:175-176)."""
    p, s = encode_sym(W, fmt)
    return np.ascontiguousarray(p.T), np.ascontiguousarray(s.T)


def take_decode_sym(packed_rm: np.ndarray, scale_rm: np.ndarray, ids: np.ndarray, K: int, fmt: Fmt):
    dt = fmt.np_dtype
    j = np.arange(K)
    word = packed_rm[ids][:, j // fmt.npu]
    shift = ((j % fmt.npu) * fmt.nbit).astype(np.uint32)[None, :]
    q = (word >> shift) & np.uint32((1 << fmt.nbit) - 1)
    w = (q.astype(dt) - dt(fmt.maxint)).astype(dt)
    return (w * scale_rm[ids][:, j // fmt.group]).astype(dt)


# --------------------------------------------------------------------------------------
# variant schemes: uint16-storage int3 and fp4  (same te_encode_sym / te_decode_sym, other branches)
# --------------------------------------------------------------------------------------


def f_to_f4(v: np.ndarray, dt) -> np.ndarray:
    """_tir_f32_to_uint_to_f4 / _tir_f16_to_uint_to_f4 (quantization.py:51-75): sign | 3-bit exponent code.
    e > bias: min(e - bias + M_h, 7) with M_h the top mantissa bit (round half up on the exponent); e == bias: 1;
    below: 0.  bias = 120 (f32) / 8 (f16)."""
    if dt == np.float32:
        u = np.ascontiguousarray(v, np.float32).view(np.uint32).astype(np.int64)
        m_h, e, sg, bias = (u >> 22) & 1, (u >> 23) & 255, u >> 31, 120
    else:
        u = np.ascontiguousarray(v, np.float16).view(np.uint16).astype(np.int64)
        m_h, e, sg, bias = (u >> 9) & 1, (u >> 10) & 31, u >> 15, 8
    e4 = np.where(e > bias, np.minimum(e - bias + m_h, 7), np.where(e == bias, 1, 0))
    return ((sg << 3) | e4).astype(np.uint32)


def f4_to_f(code: np.ndarray, dt) -> np.ndarray:
    """_tir_u32_to_f4_to_f32 / _f16 (quantization.py:78-101): 0 for exponent code 0, else sign * 2^(code - 7)."""
    sg, e4 = (code >> 3) & 1, code & 7
    if dt == np.float32:
        bits = (((e4 | 120) | (sg << 8)).astype(np.uint32) << np.uint32(23)).view(np.float32)
    else:
        bits = (((e4 | 8) | (sg << 5)).astype(np.uint16) << np.uint16(10)).view(np.float16)
    return np.where(e4 == 0, dt(0), bits).astype(dt)


def encode_variant(W: np.ndarray, fmt: Fmt, transpose: bool = True):
    """te_encode_sym (quantization.py:142-177) for the variant schemes: storage words of fmt.storage_nbit bits;
    fp4: scale = max(max_abs, 1e-4) (not divided), code = f4(W / scale)."""
    dt, sdt = fmt.np_dtype, storage_dtype(fmt)
    N, K = W.shape
    g, npu, maxint = fmt.group, fmt.npu, fmt.maxint
    G, R = fmt.n_groups(K), fmt.n_words(K)
    Wp = np.full((N, G * g), np.finfo(dt).min, dtype=dt)
    Wp[:, :K] = np.abs(W)
    mx = np.maximum(Wp.reshape(N, G, g).max(axis=2), dt(1e-4))
    scale = mx.astype(dt) if is_fp4(fmt) else (mx / dt(maxint)).astype(dt)
    t = (W / np.repeat(scale, g, axis=1)[:, :K]).astype(dt)
    if is_fp4(fmt):
        q = f_to_f4(t, dt)
    else:
        q = np.minimum(np.maximum(_round_half_away((t + dt(maxint)).astype(dt)), 0.0), float(2 * maxint)).astype(np.uint32)
    qp = np.zeros((N, R * npu), dtype=np.uint32)
    qp[:, :K] = q
    qp = qp.reshape(N, R, npu)
    packed = np.zeros((N, R), dtype=np.uint32)
    for tpos in range(npu):
        packed |= qp[:, :, tpos] << np.uint32(tpos * fmt.nbit)
    packed = packed.astype(sdt)
    return (np.ascontiguousarray(packed.T), np.ascontiguousarray(scale.T)) if transpose else (packed, scale)


def decode_variant(packed: np.ndarray, scale: np.ndarray, K: int, fmt: Fmt, transpose_output: bool = False):
    dt = fmt.np_dtype
    k = np.arange(K)
    q = (packed[k // fmt.npu, :].astype(np.uint32) >> ((k % fmt.npu) * fmt.nbit).astype(np.uint32)[:, None]) & np.uint32((1 << fmt.nbit) - 1)
    v = f4_to_f(q, dt) if is_fp4(fmt) else (q.astype(dt) - dt(fmt.maxint)).astype(dt)
    w = (v * scale[k // fmt.group, :]).astype(dt)
    return np.ascontiguousarray(w.T) if transpose_output else w


def take_decode_variant(packed_rm, scale_rm, ids, K: int, fmt: Fmt):
    dt = fmt.np_dtype
    j = np.arange(K)
    q = (packed_rm[ids][:, j // fmt.npu].astype(np.uint32) >> ((j % fmt.npu) * fmt.nbit).astype(np.uint32)[None, :]) & np.uint32((1 << fmt.nbit) - 1)
    v = f4_to_f(q, dt) if is_fp4(fmt) else (q.astype(dt) - dt(fmt.maxint)).astype(dt)
    return (v * scale_rm[ids][:, j // fmt.group]).astype(dt)


def gemv_f64_variant(packed, scale, x, fmt: Fmt, K=None):
    K = x.size if K is None else K
    w = decode_variant(packed, scale, K, fmt).astype(np.float64)
    xv = x.reshape(-1).astype(np.float64)[:K]
    return xv @ w, np.abs(xv) @ np.abs(w)


# --------------------------------------------------------------------------------------
# asymmetric scheme  (quantization.py:105-140 te_encode_asym, :183-211 te_decode_asym, :235-255 take)
# --------------------------------------------------------------------------------------


def f32x2_to_bf16x2(v0: np.ndarray, v1: np.ndarray) -> np.ndarray:
    """_tir_f32x2_to_bf16x2_to_u32 (quantization.py:15-25), round_to_even=True: v0 in the low half."""
    res = []
    for v in (v0, v1):
        u = np.ascontiguousarray(v, dtype=np.float32).view(np.uint32).astype(np.uint64)
        bias = ((u >> np.uint64(16)) & np.uint64(1)) + np.uint64(0x7FFF)
        u = (u + bias) & np.uint64(0xFFFFFFFF)  # uint32 arithmetic wraps
        res.append(((u >> np.uint64(16)) & np.uint64(0xFFFF)).astype(np.uint32))
    return res[0] | (res[1] << np.uint32(16))


def bf16x2_to_f32x2(x: np.ndarray):
    """_tir_u32_to_bf16x2_to_f32x2 (quantization.py:28-32)."""
    lo = ((x & np.uint32(0xFFFF)) << np.uint32(16)).view(np.float32)
    hi = (((x >> np.uint32(16)) & np.uint32(0xFFFF)) << np.uint32(16)).view(np.float32)
    return lo, hi


def encode_asym(W: np.ndarray, fmt: Fmt, transpose: bool = True):
    """te_encode_asym: W[N, K] -> packed uint32[K/8, N] and
        float16: scale f16[G, N], min f16[G, N]            (three tensors, quantization.py:134-138)
        float32: scale_min uint32[G, N] = bf16(scale) | bf16(min) << 16   (:126-131; the weights are quantized
                 with the UNROUNDED f32 scale / min, the decoder only ever sees the bf16 values)
    K must be a multiple of the group size (:107 assert).  transpose=False: [N, K/8] / [N, G] (embedding)."""
    dt = fmt.np_dtype
    assert not fmt.sym and W.dtype == dt and W.ndim == 2
    N, K = W.shape
    g, npu = fmt.group, fmt.npu
    assert K % g == 0
    G = K // g
    qmax = (1 << fmt.nbit) - 1
    Wg = W.reshape(N, G, g)
    mn, mx = Wg.min(axis=2), Wg.max(axis=2)
    with np.errstate(divide="ignore", invalid="ignore"):
        scale = ((mx - mn).astype(dt) / dt(qmax)).astype(dt)
        t = ((Wg - mn[:, :, None]).astype(dt) / scale[:, :, None]).astype(dt)
    q = _round_half_away(t)
    q = np.nan_to_num(q, nan=0.0, posinf=float(qmax), neginf=0.0)  # a constant group divides by zero: UB in the reference
    q = np.minimum(np.maximum(q, 0.0), float(qmax)).astype(np.uint32).reshape(N, K // npu, npu)
    packed = np.zeros((N, K // npu), dtype=np.uint32)
    for tpos in range(npu):
        packed |= q[:, :, tpos] << np.uint32(tpos * fmt.nbit)
    if fmt.dtype == "float32":
        sm = f32x2_to_bf16x2(scale, mn)
        return (np.ascontiguousarray(packed.T), np.ascontiguousarray(sm.T)) if transpose else (packed, sm)
    if transpose:
        return np.ascontiguousarray(packed.T), np.ascontiguousarray(scale.T), np.ascontiguousarray(mn.T)
    return packed, scale, mn


def _asym_scale_min(fmt: Fmt, scale, mn):
    if fmt.dtype == "float32":
        return bf16x2_to_f32x2(scale)
    return scale, mn


def decode_asym(packed: np.ndarray, scale: np.ndarray, mn, K: int, fmt: Fmt, transpose_output: bool = False):
    """te_decode_asym: w[k, n] = dtype(q) * scale + min, the product and the sum each rounded to dtype (the TVM C
    and LLVM CPU builds of the reference do not contract them into an fma; quantization.py:196)."""
    dt = fmt.np_dtype
    k = np.arange(K)
    q = (packed[k // fmt.npu, :] >> ((k % fmt.npu) * fmt.nbit).astype(np.uint32)[:, None]) & np.uint32((1 << fmt.nbit) - 1)
    s, m = _asym_scale_min(fmt, scale, mn)
    w = (q.astype(dt) * s[k // fmt.group, :]).astype(dt)
    w = (w + m[k // fmt.group, :]).astype(dt)
    return np.ascontiguousarray(w.T) if transpose_output else w


def take_decode_asym(packed_rm, scale_rm, mn_rm, ids, K: int, fmt: Fmt):
    """te_take_decode_asym (quantization.py:235-255): row-major (V, K/8) table."""
    dt = fmt.np_dtype
    j = np.arange(K)
    q = (packed_rm[ids][:, j // fmt.npu] >> ((j % fmt.npu) * fmt.nbit).astype(np.uint32)[None, :]) & np.uint32((1 << fmt.nbit) - 1)
    s, m = _asym_scale_min(fmt, scale_rm, mn_rm)
    w = (q.astype(dt) * s[ids][:, j // fmt.group]).astype(dt)
    return (w + m[ids][:, j // fmt.group]).astype(dt)


def gemv_ref_asym(packed, scale, mn, x, fmt: Fmt, epilogue=EP_NONE, ep0=None):
    """fused_decode4_matmul5_before / _fp16_before (dispatch/llama/main.py:4397-4415, 4979-4997): dtype accumulate."""
    dt = fmt.np_dtype
    K = x.size
    x = x.reshape(-1).astype(dt)
    w = decode_asym(packed, scale, mn, K, fmt)
    acc = np.zeros(packed.shape[1], dtype=dt)
    for k in range(K):
        acc = (acc + (x[k] * w[k]).astype(dt)).astype(dt)
    return apply_epilogue(acc, epilogue, ep0)


def gemv_f64_asym(packed, scale, mn, x, fmt: Fmt):
    w = decode_asym(packed, scale, mn, x.size, fmt).astype(np.float64)
    xv = x.reshape(-1).astype(np.float64)
    return xv @ w, np.abs(xv) @ np.abs(w)


# --------------------------------------------------------------------------------------
# epilogues, op-for-op in dtype like the fused PrimFuncs
# --------------------------------------------------------------------------------------


def _erf(x: np.ndarray) -> np.ndarray:
    return np.vectorize(math.erf, otypes=[np.float64])(x.astype(np.float64))


def apply_epilogue(y: np.ndarray, epilogue: int, ep0=None, ep1=None, ep2=None):
    """y: matmul result already rounded to dtype.  Follows
    add       dispatch_tir_operator_adreno.py:86-91   out = r + y
    multiply  :150-152                                out = other * y
    silu      :216-228                                out = y * sigmoid(y)
    cast_f32  :364-369                                out = f32(y)
    bias      gpt_neox/dolly_v2_3b_mod.py:41-46       out = y + bias[n]
    bias_gelu :204-251   v=y+b; t=v*f16(0.7071..); e=f16(erf(f32(t))); out=v*(f16(.5)+e*f16(.5))
    bias_add2 :88-110    out = ((y + bias) + r1) + r2   (fused_NT_matmul1_add3_add5_add5)
    """
    dt = y.dtype.type
    if epilogue == EP_NONE:
        return y
    if epilogue == EP_ADD:
        return (ep0.astype(dt) + y).astype(dt)
    if epilogue == EP_MULTIPLY:
        return (ep0.astype(dt) * y).astype(dt)
    if epilogue == EP_SILU:
        # TVM legalises sigmoid(y) to 1/(1+exp(-y)) with every op typed in dtype (exp via expf)
        e = np.exp((-y).astype(np.float32)).astype(dt)
        sig = (dt(1) / (dt(1) + e).astype(dt)).astype(dt)
        return (y * sig).astype(dt)
    if epilogue == EP_CAST_F32:
        return y.astype(np.float32)
    if epilogue == EP_BIAS:
        return (y + ep0.astype(dt)).astype(dt)
    if epilogue == EP_BIAS_GELU:
        v = (y + ep0.astype(dt)).astype(dt)
        t = (v * dt(0.70710678118654757)).astype(dt)
        e = _erf(t.astype(np.float32)).astype(np.float32).astype(dt)
        h = (dt(0.5) + (e * dt(0.5)).astype(dt)).astype(dt)
        return (v * h).astype(dt)
    if epilogue == EP_BIAS_ADD2:
        v = (y + ep0.astype(dt)).astype(dt)
        v = (v + ep1.astype(dt)).astype(dt)
        return (v + ep2.astype(dt)).astype(dt)
    raise ValueError(epilogue)


# --------------------------------------------------------------------------------------
# a5: fused decode GEMV, reference semantics  (dispatch_tir_operator_adreno.py:5-23 et al.)
# --------------------------------------------------------------------------------------


def gemv_ref(packed, scale, x, fmt: Fmt, epilogue=EP_NONE, ep0=None, K=None, reassociate=False):
    """y[n] = sum_k x[k]*w[k,n], accumulated IN dtype, init 0, k ascending (the reference's chain).

    reassociate=True mirrors the int3 "after" kernels ((x*(q-3))*s, dispatch/llama/main.py:5547).
    Vectorised over n; sequential over k (so O(K) numpy steps: keep K*N modest).
    """
    dt = fmt.np_dtype
    K = x.size if K is None else K
    x = x.reshape(-1).astype(dt)
    N = packed.shape[1]
    acc = np.zeros(N, dtype=dt)
    mask = np.uint32((1 << fmt.nbit) - 1)
    for k in range(K):
        q = (packed[k // fmt.npu] >> np.uint32((k % fmt.npu) * fmt.nbit)) & mask
        qf = (q.astype(dt) - dt(fmt.maxint)).astype(dt)
        s = scale[k // fmt.group]
        if reassociate:
            term = ((x[k] * qf).astype(dt) * s).astype(dt)
        else:
            term = (x[k] * (qf * s).astype(dt)).astype(dt)
        acc = (acc + term).astype(dt)
    return apply_epilogue(acc, epilogue, ep0)


def gemv_f64(packed, scale, x, fmt: Fmt, K=None):
    """f64 evaluation of the same dequantised weights (the 'truth' tolerances are stated against)
    plus sum_k |x_k w_kn| for the error bound."""
    K = x.size if K is None else K
    w = decode_sym(packed, scale, K, fmt).astype(np.float64)
    xv = x.reshape(-1).astype(np.float64)[:K]
    return xv @ w, np.abs(xv) @ np.abs(w)


def gemv_f64_unrounded(packed, scale, x, fmt: Fmt, K=None, chunk: int = 2048):
    """f64 evaluation with the UNROUNDED weights (q - maxint) * s (s = the stored scale, exact in f64): what the
    format means in real arithmetic, before the reference rounds every weight to dtype (quantization.py:222-224).
    Returns (y, sum_k |x_k w_kn|, sum_k off_k |x_k| s_kn) where off_k is the integer offset the tcgen05 decode
    program carries through its tensor-core sum for element k (4-bit: 1024 + maxint for even nibbles of a word,
    64 + maxint for odd ones; 8-bit: 1024 + maxint; csrc/gemv_tc.cuh) - the magnitude that cancels in  T - C  and bounds that kernel's extra error."""
    K = x.size if K is None else K
    xv = x.reshape(-1).astype(np.float64)[:K]
    N = packed.shape[1]
    k = np.arange(K)
    shift = ((k % fmt.npu) * fmt.nbit).astype(np.uint32)[:, None]
    off = np.where((k % fmt.npu) % 2 == 0, 1024.0 + fmt.maxint, 64.0 + fmt.maxint)
    if fmt.nbit == 8:  # q8f16_0 on the tcgen05 program: every byte sits under the exponent of 1024
        off = np.full(K, 1024.0 + fmt.maxint)
    y = np.empty(N, np.float64)
    ab = np.empty(N, np.float64)
    cc = np.empty(N, np.float64)
    for n0 in range(0, N, chunk):
        sl = slice(n0, min(N, n0 + chunk))
        q = (packed[k // fmt.npu, sl] >> shift) & np.uint32((1 << fmt.nbit) - 1)
        sc = scale[k // fmt.group, sl].astype(np.float64)
        w = (q.astype(np.float64) - fmt.maxint) * sc
        y[sl] = xv @ w
        ab[sl] = np.abs(xv) @ np.abs(w)
        cc[sl] = (np.abs(xv) * off) @ np.abs(sc)
        if fmt.nbit == 4:
            # int8 formulation of the 4-bit program (csrc/gemv_tc.cuh, kI8): a group's activations become integers at
            # 2^-21 of the group's largest magnitude -> error <= 2^-21 amax_g sum_{k in g} |w_k|
            ng = -(-K // fmt.group)
            xa = np.zeros(ng * fmt.group)
            xa[:K] = np.abs(xv)
            amax = np.repeat(xa.reshape(ng, fmt.group).max(axis=1), fmt.group)[:K]
            cc[sl] += 2.0 * (amax @ np.abs(w))
    return y, ab, cc


def tolerance_group_scaled(y64u, abs_sum_u, cancel_sum, out_f32: bool) -> np.ndarray:
    """Bound of the tcgen05 decode program against the UNROUNDED truth (gemv_f64_unrounded):
        |y - y64u| <= rel*|y64u| + 2^-14 * sum|x w| + 2^-22 * (sum_k off_k |x_k| s + 2 sum_g amax_g sum_{k in g} |w_k|) + 2^-24
    the SURVEY 8c(ii) bound plus the cancellation term: T = sum (off + q) x and C = sum (off + maxint) x are each
    rounded to f32 (relative 2^-24 per accumulation step of a magnitude <= sum off |x|) before they are
    subtracted.  For zero-mean activations the third term is about as large as the second; it dominates when a
    group's activations share a sign (large mean, outlier channels) or when |q - maxint| is small."""
    rel = 2.0**-22 if out_f32 else 2.0**-10
    return rel * np.abs(y64u) + 2.0**-14 * abs_sum_u + 2.0**-22 * cancel_sum + 2.0**-24


# --------------------------------------------------------------------------------------
# a7: prefill dense NT GEMM  (transform/transpose_matmul.py:34-101; main.py:3884-3898)
# --------------------------------------------------------------------------------------


def nt_matmul_ref(X, T, dtype=np.float16):
    """Y[m,n] = sum_k X[m,k]*T[n,k] accumulated in dtype, k ascending (reference semantics)."""
    M, K = X.shape
    acc = np.zeros((M, T.shape[0]), dtype=dtype)
    Tt = T.T
    for k in range(K):
        acc = (acc + (X[:, k : k + 1] * Tt[k : k + 1, :]).astype(dtype)).astype(dtype)
    return acc


def gemm_f64(packed, scale, X, fmt: Fmt):
    K = X.shape[-1]
    w = decode_sym(packed, scale, K, fmt).astype(np.float64)
    X64 = X.astype(np.float64)
    return X64 @ w, np.abs(X64) @ np.abs(w)


# --------------------------------------------------------------------------------------
# tolerances  
# --------------------------------------------------------------------------------------


def tolerance(y64: np.ndarray, abs_sum: np.ndarray, out_f32: bool, group_scaled: bool = False) -> np.ndarray:
    """|y - y64| <= rel*|y64| + 2^-14 * sum|x w| + 2^-24 ; rel = 2^-10 (f16 out) or 2^-22 (f32 out).

    y64 is the f64 evaluation on the reference's dequantized weights f16((q - maxint) * s).
    group_scaled: the kernel applies the group scale once to the f32 group sum  s * sum (q - maxint) x
    (the tcgen05 decode program) instead of rounding every weight to f16 first; each reference weight is
    within 2^-11 relative of (q - maxint) * s, so the bound gains a deterministic 2^-11 * sum|x w|
    (the measured deviation is ~sqrt(K) times smaller, and the result is closer to exact arithmetic)."""
    rel = 2.0**-22 if out_f32 else 2.0**-10
    tol = rel * np.abs(y64) + 2.0**-14 * abs_sum + 2.0**-24
    if group_scaled:
        tol = tol + 2.0**-11 * abs_sum
    return tol


# --------------------------------------------------------------------------------------
# algorithmic bytes  
# --------------------------------------------------------------------------------------


def gemv_bytes(K: int, N: int, fmt: Fmt, epilogue=EP_NONE) -> int:
    sb = 2 if fmt.dtype == "float16" else 4
    ob = 4 if (epilogue == EP_CAST_F32 or fmt.dtype == "float32") else 2
    b = (fmt.storage_nbit // 8) * fmt.n_words(K) * N + sb * fmt.n_groups(K) * N + sb * K + ob * N
    if not fmt.sym and fmt.dtype == "float16":
        b += sb * fmt.n_groups(K) * N  # the min tensor
    if epilogue in (EP_ADD, EP_MULTIPLY):
        b += ob * N
    if epilogue in (EP_BIAS, EP_BIAS_GELU):
        b += sb * N
    return b


# --------------------------------------------------------------------------------------
# synthetic data  (seed 0xB200, W ~ N(0, 0.02^2), x ~ N(0,1))
# --------------------------------------------------------------------------------------

SEED = 0xB200


def synth_weight(N: int, K: int, fmt: Fmt, seed: int = SEED):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((N, K), dtype=np.float32) * 0.02).astype(fmt.np_dtype)


def synth_act(shape, fmt: Fmt, seed: int = SEED + 1, std: float = 1.0):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(shape, dtype=np.float32) * std).astype(fmt.np_dtype)
