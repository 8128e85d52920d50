"""uint16-storage int3 and fp4 on the GPU (csrc/variants.cu): encode / dequantize / take_decode bit-exact against the
reference-run goldens (tests/golden/variants.npz) and the oracle, the decode-step GEMV inside the stated tolerance
with op-for-op epilogues."""
import os

import numpy as np
import pytest
import torch

from mlc_llm_b200 import dispatch, ops
from oracle import quant_oracle as qo
from util import beq, host

pytestmark = pytest.mark.gpu
NAMES = list(qo.VARIANT_FORMATS)


def dv(a):
    if a.dtype == np.uint32:
        a = a.view(np.int32)
    if a.dtype == np.uint16:
        a = a.view(np.int16)
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def hv(t, like):
    a = t.detach().cpu().numpy()
    return a.view(like) if a.dtype in (np.int32, np.int16) else a


@pytest.fixture(scope="module")
def var(golden_dir):
    return np.load(os.path.join(golden_dir, "variants.npz"))


def test_against_reference_goldens(var):
    for c in sorted({k.split("/")[0] for k in var.files}):
        name = c.split("_N")[0]
        fmt = qo.VARIANT_FORMATS[name]
        W = var[c + "/W"]
        K = W.shape[1]
        sdt = qo.storage_dtype(fmt)
        for tr, sfx in ((True, ""), (False, "_rm")):
            p, s = ops.encode(dv(W), name, transpose=tr)
            assert beq(hv(p, sdt), var[f"{c}/packed{sfx}"]), (c, tr)
            assert beq(host(s), var[f"{c}/scale{sfx}"]), (c, tr)
        if c + "/w" not in var.files:
            continue
        p, s = dv(var[c + "/packed"]), dv(var[c + "/scale"])
        assert beq(host(ops.dequantize(p, s, K, name)), var[c + "/w"]), c
        assert beq(host(ops.dequantize(p, s, K, name, True)), var[c + "/w_t"]), c
        assert beq(host(ops.dequantize(dv(var[c + "/rand_packed"]), dv(var[c + "/rand_scale"]), K, name)), var[c + "/rand_w"]), c
        if c + "/take_out" in var.files:
            out = ops.take_decode(dv(var[c + "/packed_rm"]), dv(var[c + "/scale_rm"]), dv(var[c + "/take_ids"]), K, name)
            assert beq(host(out), var[c + "/take_out"]), c


@pytest.mark.parametrize("name", NAMES)
def test_decode_every_code_and_gemv(name):
    fmt = qo.VARIANT_FORMATS[name]
    f32 = fmt.dtype == "float32"
    sdt = qo.storage_dtype(fmt)
    rng = np.random.default_rng(9)
    for K, N in ((4096, 4096), (4000 if fmt.nbit == 3 else 4064, 264), (11008 if fmt.nbit != 3 else 11000, 512)):
        packed = rng.integers(0, 2**fmt.storage_nbit, size=(fmt.n_words(K), N), dtype=np.uint64).astype(sdt)
        scale = (rng.random((fmt.n_groups(K), N), dtype=np.float32) * 0.01).astype(fmt.np_dtype)
        p, s = dv(packed), dv(scale)
        for tr in (False, True):
            assert beq(host(ops.dequantize(p, s, K, name, tr)), qo.decode_variant(packed, scale, K, fmt, tr)), (K, N, tr)
        x = qo.synth_act(K, fmt, seed=K)
        y0 = host(ops.gemv(p, s, dv(x), name)).reshape(-1)
        y64, ab = qo.gemv_f64_variant(packed, scale, x, fmt)
        err = np.abs(y0.astype(np.float64) - y64)
        assert (err <= qo.tolerance(y64, ab, f32)).all(), float((err / qo.tolerance(y64, ab, f32)).max())
        epv = qo.synth_act(N, fmt, seed=3)
        for ep in ("add", "multiply", "bias"):
            y = host(ops.gemv(p, s, dv(x), name, ep, dv(epv))).reshape(-1)
            assert beq(y, qo.apply_epilogue(y0, qo.EPILOGUES[ep], epv)), ep


def test_dispatch_and_unsupported_paths():
    name = "q3f16_u16"
    fmt = qo.VARIANT_FORMATS[name]
    K, N = 4000, 256
    assert dispatch.lookup(name, K, N, "none", m_is_one=True) is not None
    assert dispatch.lookup_decode(name, K, N, True) is not None
    assert dispatch.lookup(name, K, N, "none", m_is_one=False) is None  # prefill: decode + NT_matmul, like the reference
