"""uint16-storage int3 and fp4: the oracle against vectors produced by the reference's own
te_encode_sym / te_decode_sym / te_take_decode_sym with storage_nbit=16 / mode="fp4" (quantization.py:51-101,
142-177, 213-270) through TVM's C backend.  Bit-exact.  fp4 + float16: the reference's decoder does not build
(it reinterprets a uint32 as float16), so only its encoder is pinned; the f32 scheme pins both."""
import os

import numpy as np
import pytest

from oracle import quant_oracle as qo


def beq(a, b):
    return a.shape == b.shape and a.dtype == b.dtype and np.ascontiguousarray(a).tobytes() == np.ascontiguousarray(b).tobytes()


@pytest.fixture(scope="module")
def var(golden_dir):
    return np.load(os.path.join(golden_dir, "variants.npz"))


def cases(z):
    return sorted({k.split("/")[0] for k in z.files})


def test_encode_bit_exact(var):
    assert {c.split("_N")[0] for c in cases(var)} == set(qo.VARIANT_FORMATS)
    for c in cases(var):
        fmt = qo.VARIANT_FORMATS[c.split("_N")[0]]
        for tr, sfx in ((True, ""), (False, "_rm")):
            p, s = qo.encode_variant(var[c + "/W"], fmt, transpose=tr)
            assert beq(p, var[f"{c}/packed{sfx}"]), (c, tr)
            assert beq(s, var[f"{c}/scale{sfx}"]), (c, tr)


def test_decode_and_take_bit_exact(var):
    n = 0
    for c in cases(var):
        if c + "/w" not in var.files:
            continue
        fmt = qo.VARIANT_FORMATS[c.split("_N")[0]]
        K = var[c + "/W"].shape[1]
        assert beq(qo.decode_variant(var[c + "/packed"], var[c + "/scale"], K, fmt), var[c + "/w"]), c
        assert beq(qo.decode_variant(var[c + "/packed"], var[c + "/scale"], K, fmt, True), var[c + "/w_t"]), c
        assert beq(qo.decode_variant(var[c + "/rand_packed"], var[c + "/rand_scale"], K, fmt), var[c + "/rand_w"]), c
        if c + "/take_out" in var.files:
            out = qo.take_decode_variant(var[c + "/packed_rm"], var[c + "/scale_rm"], var[c + "/take_ids"], K, fmt)
            assert beq(out, var[c + "/take_out"]), c
        n += 1
    assert n >= 4


def test_fp4_code_table():
    """Every fp4 code decodes to 0 or +-2^(e-7) (quantization.py:78-101) and re-encodes to itself."""
    for dt in (np.float32, np.float16):
        codes = np.arange(16, dtype=np.uint32)
        v = qo.f4_to_f(codes, dt)
        assert [float(x) for x in v[:8]] == [0.0] + [2.0**(e - 7) for e in range(1, 8)]
        assert np.array_equal(v[8:], -v[:8])
        back = qo.f_to_f4(v, dt)
        assert np.array_equal(back[1:8], codes[1:8]) and np.array_equal(back[9:], codes[9:])
