"""The prefill layer-sequence driver (mlc_llm_b200/prefill_stack.py): the GEMM chain of a 2-layer model against the
oracle evaluating the same chain of reference operators (decode* + NT_matmul* + fused epilogues,
mlc_llm/dispatch/gpt_neox/dolly_v2_3b_mod.py:26-275, dispatch/llama/main.py:2832-3660) on the same packed weights."""
import numpy as np
import pytest
import torch

from mlc_llm_b200.decode_stack import DecodeStack, ModelShape
from mlc_llm_b200.prefill_stack import PrefillStack
from oracle import quant_oracle as qo
from util import host

pytestmark = pytest.mark.gpu


def _lin(w, X, fmt, ep=qo.EP_NONE, e0=None, e1=None, e2=None):
    y64, _ = qo.gemm_f64(host(w.packed), host(w.scale), X, fmt)
    return qo.apply_epilogue(y64.astype(np.float16), ep, e0, e1, e2)


@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0"])
def test_neox_prefill_chain(name):
    fmt = qo.FORMATS[name]
    shape = ModelShape("tiny-neox", "gpt_neox", 640, 2560, 2, 1000)
    st = DecodeStack(shape, name, layers=2)
    M = 300
    pf = PrefillStack(st, M)
    X0 = (np.random.default_rng(0).standard_normal((M, shape.hidden)) * 0.5).astype(np.float16)
    pf.x_in.copy_(torch.from_numpy(X0))
    Y = host(pf.run().clone())
    assert pf.gemms_per_pass == 12
    x = X0
    for li in range(2):
        w, b = st.layers[li], {k: np.broadcast_to(host(v), (M, v.numel())) for k, v in st.biases[li].items()}
        q = _lin(w["q"], x, fmt, qo.EP_BIAS, b["q"])
        a = _lin(w["h_to_4h"], x, fmt, qo.EP_BIAS_GELU, b["h_to_4h"])
        m = _lin(w["4h_to_h"], a, fmt, qo.EP_BIAS, b["4h_to_h"])
        x = _lin(w["dense"], q, fmt, qo.EP_BIAS_ADD2, b["dense"], m, x)
    assert np.isfinite(Y).all()
    assert np.abs(Y.astype(np.float64) - x.astype(np.float64)).max() <= 2e-2 * np.abs(x.astype(np.float64)).max()


def test_llama_prefill_chain():
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    shape = ModelShape("tiny-llama", "llama", 512, 1408, 2, 1000)
    st = DecodeStack(shape, name, layers=2)
    M = 260
    pf = PrefillStack(st, M)
    X0 = (np.random.default_rng(1).standard_normal((M, shape.hidden)) * 0.5).astype(np.float16)
    pf.x_in.copy_(torch.from_numpy(X0))
    Y = host(pf.run().clone())
    x = X0
    for li in range(2):
        w = st.layers[li]
        q = _lin(w["q"], x, fmt)
        x1 = _lin(w["o"], q, fmt, qo.EP_ADD, x)
        g = _lin(w["gate"], x1, fmt, qo.EP_SILU)
        a = _lin(w["up"], x1, fmt, qo.EP_MULTIPLY, g)
        x = _lin(w["down"], a, fmt, qo.EP_ADD, x1)
    assert np.abs(Y.astype(np.float64) - x.astype(np.float64)).max() <= 2e-2 * np.abs(x.astype(np.float64)).max()
