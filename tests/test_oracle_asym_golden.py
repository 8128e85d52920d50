"""The asymmetric-scheme oracle (numpy + C) against vectors produced by the reference's own te_encode_asym /
te_decode_asym / te_take_decode_asym (mlc_llm/transform/quantization.py:105-140,183-211,235-255) and its
unscheduled fused GEMVs fused_decode4_matmul5_before / _fp16_before (dispatch/llama/main.py:4397-4415,4979-4997),
run through TVM's C backend by oracle/gen_golden.py.  Everything is bit-exact."""
import os

import numpy as np
import pytest

from oracle import c_oracle as co
from oracle import quant_oracle as qo
from oracle.golden_inputs import sha


def beq(a, b):
    return a.shape == b.shape and a.dtype == b.dtype and np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


@pytest.fixture(scope="module")
def asym(golden_dir):
    return np.load(os.path.join(golden_dir, "asym.npz"))


def _cases(z):
    return sorted({k.split("/")[0] for k in z.files if "_N" in k.split("/")[0]})


def _aux(z, c, prefix=""):
    mn = z[f"{c}/{prefix}min"] if f"{c}/{prefix}min" in z.files else None
    return z[f"{c}/{prefix}scale"], mn


def test_cases_present(asym):
    assert {c.split("_N")[0] for c in _cases(asym)} == set(qo.ASYM_FORMATS)


def test_encode_asym_bit_exact(asym):
    for c in _cases(asym):
        fmt = qo.ASYM_FORMATS[c.split("_N")[0]]
        for tr, sfx in ((True, ""), (False, "_rm")):
            r = qo.encode_asym(asym[c + "/W"], fmt, transpose=tr)
            assert beq(r[0], asym[f"{c}/packed{sfx}"]), (c, tr)
            assert beq(r[1], asym[f"{c}/scale{sfx}"]), (c, tr)
            if fmt.dtype == "float16":
                assert beq(r[2], asym[f"{c}/min{sfx}"]), (c, tr)


def test_decode_asym_bit_exact_numpy_and_c(asym):
    for c in _cases(asym):
        fmt = qo.ASYM_FORMATS[c.split("_N")[0]]
        K = asym[c + "/W"].shape[1]
        s, m = _aux(asym, c)
        p = asym[c + "/packed"]
        for tr, key in ((False, "w"), (True, "w_t")):
            assert beq(qo.decode_asym(p, s, m, K, fmt, tr), asym[f"{c}/{key}"]), (c, key)
            assert beq(co.decode_asym(p, s, m, K, fmt, tr), asym[f"{c}/{key}"]), (c, key, "C")
        rs, rm = _aux(asym, c, "rand_")
        assert beq(qo.decode_asym(asym[c + "/rand_packed"], rs, rm, K, fmt), asym[c + "/rand_w"]), c
        assert beq(co.decode_asym(asym[c + "/rand_packed"], rs, rm, K, fmt), asym[c + "/rand_w"]), c


def test_take_decode_asym_bit_exact(asym):
    for c in _cases(asym):
        fmt = qo.ASYM_FORMATS[c.split("_N")[0]]
        K = asym[c + "/W"].shape[1]
        s, m = _aux(asym, c, "")
        srm = asym[c + "/scale_rm"]
        mrm = asym[c + "/min_rm"] if c + "/min_rm" in asym.files else None
        out = qo.take_decode_asym(asym[c + "/packed_rm"], srm, mrm, asym[c + "/take_ids"], K, fmt)
        assert beq(out, asym[c + "/take_out"]), c


def test_bf16_pack_round_to_nearest_even():
    """quantization.py:15-25: ties go to the even bf16, carries propagate into the exponent."""
    v = np.array([1.0, 1.00390625, 1.01171875, -0.0, 3.3895314e38], np.float32)  # exact, tie->even (down), tie->even (up)
    u = qo.f32x2_to_bf16x2(v, v)
    lo, hi = qo.bf16x2_to_f32x2(u)
    assert np.array_equal(lo, hi)
    assert lo[0] == 1.0 and lo[1] == 1.0 and lo[2] == np.float32(1.015625)


def test_fused_asym_gemv_bit_exact(asym):
    """The C oracle reproduces the reference's dtype-accumulating fused GEMV bit for bit (4096 x 4096)."""
    for ci, (fname, name) in enumerate([("fused_decode4_matmul5_before", "q4f32_asym"), ("fused_decode4_matmul5_fp16_before", "q4f16_asym")]):
        fmt = qo.ASYM_FORMATS[name]
        key = f"{name}/{fname}"
        packed, aux, x = synth_asym_fused(fmt, int(asym[key + "/seed"][0]))
        assert sha(packed, *aux, x) == bytes(asym[key + "/sha"]).decode()
        y = co.gemv_ref_asym(packed, aux[0], aux[1] if len(aux) > 1 else None, x, fmt)
        assert beq(y, asym[key + "/y"]), fname


def synth_asym_fused(fmt, seed, K=4096, N=4096):
    rng = np.random.default_rng(seed)
    packed = rng.integers(0, 2**32, size=(K // 8, N), dtype=np.uint64).astype(np.uint32)
    sc = rng.random((K // 32, N), dtype=np.float32) * 0.01
    mn = -rng.random((K // 32, N), dtype=np.float32) * 0.08
    aux = [qo.f32x2_to_bf16x2(sc, mn)] if fmt.dtype == "float32" else [sc.astype(np.float16), mn.astype(np.float16)]
    x = rng.standard_normal(K, dtype=np.float32).astype(fmt.np_dtype)
    return packed, aux, x
