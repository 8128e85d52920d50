"""GPU parity of the asymmetric int4 schemes (the reference CLI's default: quantization.py:105-140, 183-211, 235-255;
fused GEMVs dispatch/llama/main.py:4235-5313): encode / decode / take_decode bit-exact against the reference-run
goldens (tests/golden/asym.npz) and the oracle, GEMV / GEMM inside the stated tolerance and closer to the f64 truth
than the reference's own dtype-accumulated result."""
import os

import numpy as np
import pytest
import torch

from mlc_llm_b200 import dispatch, ops
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from oracle.golden_inputs import sha
from test_oracle_asym_golden import synth_asym_fused
from util import beq, dev, host

pytestmark = pytest.mark.gpu
NAMES = list(qo.ASYM_FORMATS)


@pytest.fixture(scope="module")
def asym(golden_dir):
    return np.load(os.path.join(golden_dir, "asym.npz"))


def _cases(z):
    return sorted({k.split("/")[0] for k in z.files if "_N" in k.split("/")[0]})


def _d(a):
    return None if a is None else dev(a)


def rand_asym(fmt, K, N, seed):
    rng = np.random.default_rng(seed)
    packed = rng.integers(0, 2**32, size=(K // 8, N), dtype=np.uint64).astype(np.uint32)
    sc = rng.random((K // 32, N), dtype=np.float32) * 0.01
    mn = -rng.random((K // 32, N), dtype=np.float32) * 0.08
    if fmt.dtype == "float32":
        return packed, qo.f32x2_to_bf16x2(sc, mn), None
    return packed, sc.astype(np.float16), mn.astype(np.float16)


def test_encode_decode_take_against_reference_goldens(asym):
    for c in _cases(asym):
        name = c.split("_N")[0]
        fmt = qo.ASYM_FORMATS[name]
        W = asym[c + "/W"]
        K = W.shape[1]
        f16 = fmt.dtype == "float16"
        for tr, sfx in ((True, ""), (False, "_rm")):
            r = ops.encode(dev(W), name, transpose=tr)
            assert beq(host(r[0]), asym[f"{c}/packed{sfx}"]), (c, tr)
            assert beq(host(r[1]), asym[f"{c}/scale{sfx}"]), (c, tr)
            if f16:
                assert beq(host(r[2]), asym[f"{c}/min{sfx}"]), (c, tr)
        p, s = asym[c + "/packed"], asym[c + "/scale"]
        m = asym[c + "/min"] if f16 else None
        assert beq(host(ops.dequantize(dev(p), dev(s), K, name, zmin=_d(m))), asym[c + "/w"]), c
        assert beq(host(ops.dequantize(dev(p), dev(s), K, name, True, zmin=_d(m))), asym[c + "/w_t"]), c
        rm = asym[c + "/rand_min"] if f16 else None
        assert beq(host(ops.dequantize(dev(asym[c + "/rand_packed"]), dev(asym[c + "/rand_scale"]), K, name, zmin=_d(rm))), asym[c + "/rand_w"]), c
        mrm = asym[c + "/min_rm"] if f16 else None
        out = ops.take_decode(dev(asym[c + "/packed_rm"]), dev(asym[c + "/scale_rm"]), dev(asym[c + "/take_ids"]), K, name, zmin_rm=_d(mrm))
        assert beq(host(out), asym[c + "/take_out"]), c


@pytest.mark.parametrize("name", NAMES)
def test_decode_every_code_large_shapes(name):
    fmt = qo.ASYM_FORMATS[name]
    for K, N in ((4096, 4096), (11008, 136), (32, 8)):
        p, s, m = rand_asym(fmt, K, N, seed=K + N)
        for tr in (False, True):
            assert beq(host(ops.dequantize(dev(p), dev(s), K, name, tr, zmin=_d(m))), co.decode_asym(p, s, m, K, fmt, tr)), (K, N, tr)


@pytest.mark.parametrize("name", NAMES)
def test_gemv_against_reference_run_fused_goldens(asym, name):
    fmt = qo.ASYM_FORMATS[name]
    fname = "fused_decode4_matmul5_before" if fmt.dtype == "float32" else "fused_decode4_matmul5_fp16_before"
    key = f"{name}/{fname}"
    packed, aux, x = synth_asym_fused(fmt, int(asym[key + "/seed"][0]))
    assert sha(packed, *aux, x) == bytes(asym[key + "/sha"]).decode()
    m = aux[1] if len(aux) > 1 else None
    y = host(ops.gemv(dev(packed), dev(aux[0]), dev(x), name, zmin=_d(m))).reshape(-1)
    y_ref = asym[key + "/y"]
    assert y.dtype == y_ref.dtype
    y64, ab = co.gemv_f64_asym(packed, aux[0], m, x, fmt)
    err = np.abs(y.astype(np.float64) - y64)
    assert (err <= qo.tolerance(y64, ab, fmt.dtype == "float32")).all(), float((err / qo.tolerance(y64, ab, fmt.dtype == "float32")).max())
    assert err.max() <= np.abs(y_ref.astype(np.float64) - y64).max()  # closer to the truth than the reference's own accumulate


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("K,N", [(4096, 4096), (4096, 11008), (11008, 4096), (4096, 32000), (2560, 2560), (64, 8), (96, 264)])
def test_gemv_tolerance_shapes_and_epilogues(name, K, N):
    fmt = qo.ASYM_FORMATS[name]
    f32 = fmt.dtype == "float32"
    p, s, m = rand_asym(fmt, K, N, seed=3 * K + N)
    x = qo.synth_act(K, fmt, seed=K)
    pd, sd, md, xd = dev(p), dev(s), _d(m), dev(x)
    y0 = host(ops.gemv(pd, sd, xd, name, zmin=md)).reshape(-1)
    y64, ab = co.gemv_f64_asym(p, s, m, x, fmt)
    err = np.abs(y0.astype(np.float64) - y64)
    assert (err <= qo.tolerance(y64, ab, f32)).all(), float((err / qo.tolerance(y64, ab, f32)).max())
    if N <= 4096:
        epv = qo.synth_act(N, fmt, seed=12)
        for ep in ("add", "multiply", "bias"):
            y = host(ops.gemv(pd, sd, xd, name, ep, dev(epv), zmin=md)).reshape(-1)
            assert beq(y, qo.apply_epilogue(y0, qo.EPILOGUES[ep], epv)), ep
        if not f32:
            y32 = host(ops.gemv(pd, sd, xd, name, "cast_f32", zmin=md)).reshape(-1)
            assert y32.dtype == np.float32 and beq(y32.astype(np.float16), y0)


def test_gemm_asym_f16_identity_and_tolerance():
    name = "q4f16_asym"
    fmt = qo.ASYM_FORMATS[name]
    for tile in ("single", "pair256", None):
        K, N = 320, 520
        p, s, m = rand_asym(fmt, K, N, seed=71)
        Y = host(ops.gemm(dev(p), dev(s), dev(np.eye(K, dtype=np.float16)), name, zmin=dev(m), tile=tile))
        assert beq(Y, co.decode_asym(p, s, m, K, fmt)), tile  # the B operand is the bit-exact decode
        for M, K, N in ((300, 640, 1000), (2048, 2560, 2560)):
            p, s, m = rand_asym(fmt, K, N, seed=M + N)
            X = qo.synth_act((M, K), fmt, seed=K)
            bias = qo.synth_act(N, fmt, seed=2)
            Y0 = host(ops.gemm(dev(p), dev(s), dev(X), name, zmin=dev(m), tile=tile))
            Yb = host(ops.gemm(dev(p), dev(s), dev(X), name, "bias", dev(bias), zmin=dev(m), tile=tile))
            assert beq(Yb, qo.apply_epilogue(Y0, qo.EP_BIAS, np.broadcast_to(bias, (M, N))))
            rows = np.random.default_rng(0).choice(M, 48, replace=False)
            w64 = co.decode_asym(p, s, m, K, fmt).astype(np.float64)
            y64 = X[rows].astype(np.float64) @ w64
            ab = np.abs(X[rows].astype(np.float64)) @ np.abs(w64)
            assert (np.abs(Y0[rows].astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all(), (tile, M, K, N)


def test_dispatch_asym_entries():
    """lookup(...) for the asymmetric schemes: the PrimFunc argument order (packed, scale, min, x[, operand], out)
    of fused_decode4_matmul5_fp16_* (dispatch/llama/main.py:4979) / (packed, scale_min, x, out) for f32 (:4397)."""
    K, N = 1024, 512
    for name in NAMES:
        fmt = qo.ASYM_FORMATS[name]
        p, s, m = rand_asym(fmt, K, N, seed=5)
        x = qo.synth_act(K, fmt, seed=6)
        op = dispatch.lookup(name, K, N, "none", m_is_one=True)
        assert op is not None
        out = torch.zeros(N, dtype=torch.float32 if fmt.dtype == "float32" else torch.float16, device="cuda")
        args = [dev(p), dev(s)] + ([dev(m)] if m is not None else []) + [dev(x), out]
        op(*args)
        y64, ab = co.gemv_f64_asym(p, s, m, x, fmt)
        assert (np.abs(host(out).astype(np.float64) - y64) <= qo.tolerance(y64, ab, fmt.dtype == "float32")).all()
        dop = dispatch.lookup_decode(name, K, N, True)
        wt = torch.zeros((N, K), dtype=out.dtype, device="cuda")
        dop(*([dev(p), dev(s)] + ([dev(m)] if m is not None else []) + [wt]))
        assert beq(host(wt), co.decode_asym(p, s, m, K, fmt, True))
    assert dispatch.lookup("q4f32_asym", K, N, "none", m_is_one=False) is None  # f32 prefill: a miss
    assert dispatch.lookup("q4f16_asym", K, N, "none", m_is_one=False) is not None
