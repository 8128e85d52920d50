"""The oracle (numpy + C restatements) against vectors produced by the reference's own code.

tests/golden/*.npz were written by oracle/gen_golden.py, which runs the reference's unmodified
encoding_func / decoding_func / decoding_after_taking_func (mlc_llm/transform/quantization.py) and its
unscheduled fused_decode*_matmul* PrimFuncs (dispatch_tir_operator_adreno.py:5-411,
dispatch/llama/main.py:5316-5758) through TVM's C backend.  Everything here is bit-exact.
"""
import os

import numpy as np
import pytest

from oracle import c_oracle as co
from oracle import quant_oracle as qo
from oracle.golden_inputs import epilogue_of, sha, synth_fused_inputs


def beq(a, b):
    return a.shape == b.shape and a.dtype == b.dtype and np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


@pytest.fixture(scope="module")
def encdec(golden_dir):
    return np.load(os.path.join(golden_dir, "encode_decode.npz"))


@pytest.fixture(scope="module")
def fused(golden_dir):
    return np.load(os.path.join(golden_dir, "fused_gemv.npz"))


def _cases(z):
    return sorted({k.split("/")[0] for k in z.files})


def test_case_coverage(encdec):
    names = {c.split("_N")[0] for c in _cases(encdec)}
    assert names == set(qo.FORMATS)


def test_encode_bit_exact(encdec):
    for c in _cases(encdec):
        fmt = qo.FORMATS[c.split("_N")[0]]
        p, s = qo.encode_sym(encdec[c + "/W"], fmt)
        assert beq(p, encdec[c + "/packed"]), c
        assert beq(s, encdec[c + "/scale"]), c


def test_decode_bit_exact_numpy_and_c(encdec):
    for c in _cases(encdec):
        fmt = qo.FORMATS[c.split("_N")[0]]
        K = encdec[c + "/W"].shape[1]
        for pk, sk, wk in [("packed", "scale", "w"), ("rand_packed", "rand_scale", "rand_w")]:
            p, s = encdec[f"{c}/{pk}"], encdec[f"{c}/{sk}"]
            assert beq(qo.decode_sym(p, s, K, fmt), encdec[f"{c}/{wk}"]), (c, wk)
            assert beq(co.decode(p, s, K, fmt), encdec[f"{c}/{wk}"]), (c, wk, "C")
        assert beq(qo.decode_sym(encdec[c + "/packed"], encdec[c + "/scale"], K, fmt, True), encdec[c + "/w_t"]), c
        assert beq(co.decode(encdec[c + "/packed"], encdec[c + "/scale"], K, fmt, True), encdec[c + "/w_t"]), c


def test_take_decode_bit_exact(encdec):
    n = 0
    for c in _cases(encdec):
        if c + "/take_out" not in encdec.files:
            continue
        fmt = qo.FORMATS[c.split("_N")[0]]
        W = encdec[c + "/W"]
        prm, srm = qo.encode_sym_rowmajor(W, fmt)
        assert beq(prm, encdec[c + "/packed_rm"]) and beq(srm, encdec[c + "/scale_rm"]), c
        out = qo.take_decode_sym(prm, srm, encdec[c + "/take_ids"], W.shape[1], fmt)
        assert beq(out, encdec[c + "/take_out"]), c
        n += 1
    assert n >= 4


def test_decode_all_codes_and_offsets():
    """offsets 7 / 3 / 127 and every code value (quantization.py:40-48)."""
    for fmt in qo.FORMATS.values():
        ncode = 1 << fmt.nbit
        K = fmt.group
        R, G = fmt.n_words(K), 1
        packed = np.zeros((R, ncode), np.uint32)
        for code in range(ncode):
            for k in range(K):
                packed[k // fmt.npu, code] |= np.uint32(code << ((k % fmt.npu) * fmt.nbit))
        scale = np.full((G, ncode), 0.5, fmt.np_dtype)
        w = qo.decode_sym(packed, scale, K, fmt)
        expect = (np.arange(ncode) - fmt.maxint) * 0.5
        assert np.array_equal(w.astype(np.float64), np.broadcast_to(expect, (K, ncode)))


def test_encode_decode_roundtrip_error_bound():
    for fmt in qo.FORMATS.values():
        W = qo.synth_weight(16, 8 * fmt.group, fmt, seed=5)
        p, s = qo.encode_sym(W, fmt)
        w = qo.decode_sym(p, s, W.shape[1], fmt, True)
        sfull = np.repeat(s.T.astype(np.float64), fmt.group, axis=1)
        err = np.abs(w.astype(np.float64) - W.astype(np.float64))
        # half a step, plus the dtype rounding of W/scale + maxint (up to 2*maxint ulps-of-1 in f16)
        slack = 0.5 + 0.01 + (2 * fmt.maxint) * (2.0**-10 if fmt.dtype == "float16" else 2.0**-23)
        assert (err <= slack * sfull + 1e-7).all(), fmt.name


def test_fused_gemv_c_oracle_equals_reference(fused):
    names = sorted({k.rsplit("/", 1)[0] for k in fused.files})
    assert len(names) == 12
    for n in names:
        fmt = qo.FORMATS[n.split("/")[0]]
        K, N, seed, _ = [int(v) for v in fused[n + "/meta"]]
        packed, scale, x, epv = synth_fused_inputs(fmt, K, N, seed)
        assert sha(packed, scale, x, epv) == bytes(fused[n + "/sha"]).decode(), "RNG drift: regenerate goldens"
        ep = epilogue_of(n)
        y = co.gemv_ref(packed, scale, x, fmt, ep, epv if ep in (qo.EP_ADD, qo.EP_MULTIPLY) else None)
        assert beq(y, fused[n + "/y"]), n


def test_numpy_gemv_ref_equals_c(fused):
    fmt = qo.FORMATS["q4f16_0"]
    packed, scale, x, epv = synth_fused_inputs(fmt, 256, 64, 3)
    for ep in (qo.EP_NONE, qo.EP_ADD, qo.EP_MULTIPLY, qo.EP_SILU, qo.EP_CAST_F32, qo.EP_BIAS, qo.EP_BIAS_GELU):
        a = qo.gemv_ref(packed, scale, x, fmt, ep, epv)
        b = co.gemv_ref(packed, scale, x, fmt, ep, epv)
        assert beq(a, b), ep
    for name in ("q3f16_0", "q4f32_0", "q8f16_0"):
        f = qo.FORMATS[name]
        packed, scale, x, epv = synth_fused_inputs(f, 4 * f.group + (10 if f.nbit == 3 else 0), 32, 4)
        for re in (False, True):
            assert beq(qo.gemv_ref(packed, scale, x, f, reassociate=re), co.gemv_ref(packed, scale, x, f, reassociate=re)), name
        y64, ab = qo.gemv_f64(packed, scale, x, f)
        y64c, abc = co.gemv_f64(packed, scale, x, f)
        assert np.allclose(y64, y64c, rtol=1e-12, atol=1e-12) and np.allclose(ab, abc, rtol=1e-12)


def test_reference_f16_chain_error_is_large(fused):
    """the reference's f16 chain is ~1e-1 off the f64 truth at K=4096; tolerances
    are therefore stated against f64, not against the reference's rounding."""
    n = "q4f16_0/fused_decode4_matmul3"
    fmt = qo.FORMATS["q4f16_0"]
    K, N, seed, _ = [int(v) for v in fused[n + "/meta"]]
    packed, scale, x, _ = synth_fused_inputs(fmt, K, N, seed)
    y64, _ = co.gemv_f64(packed, scale, x, fmt)
    err = np.abs(fused[n + "/y"].astype(np.float64) - y64).max()
    assert 0.01 < err < 0.5


def test_algorithmic_bytes():
    f = qo.FORMATS["q4f16_0"]
    assert abs(qo.gemv_bytes(4096, 4096, f) - 9.454e6) < 5e3
    assert abs(qo.gemv_bytes(4096, 11008, f) - 25.393e6) < 3e4
