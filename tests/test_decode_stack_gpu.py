"""GPU test of the layer-sequence driver: the graph-captured, PDL-chained, fused-launch decode step
equals the oracle evaluating the same chain of reference operators on the same packed weights."""
import numpy as np
import pytest
import torch

from mlc_llm_b200.decode_stack import DecodeStack, ModelShape
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from util import host

pytestmark = pytest.mark.gpu


def _oracle_linear(w, x, fmt, ep=qo.EP_NONE, ep0=None):
    """f64 matmul of the dequantized weights, rounded once to the model dtype, then the reference's
    dtype-typed epilogue (what the CUDA path computes up to f32 summation noise)."""
    packed, scale = host(w.packed), host(w.scale)
    y64, _ = co.gemv_f64(packed, scale, x, fmt, w.K)
    if ep == qo.EP_CAST_F32:
        return y64.astype(np.float32)
    return qo.apply_epilogue(y64.astype(fmt.np_dtype), ep, ep0)


@pytest.mark.parametrize("fuse", [True, False])
@pytest.mark.parametrize("name", ["q4f16_0", "q8f16_0", "q3f16_0"])
def test_llama_stack_matches_oracle_chain(name, fuse):
    fmt = qo.FORMATS[name]
    shape = ModelShape("tiny-llama", "llama", 640, 1760, 2, 1000)
    st = DecodeStack(shape, name, layers=2, fuse_qkv=fuse, fuse_gate_up=fuse)
    x0 = (np.random.default_rng(0).standard_normal(shape.hidden) * 0.5).astype(np.float16)
    st.x_in.copy_(torch.from_numpy(x0))
    st.capture()
    logits = host(st.decode_step().clone())
    torch.cuda.synchronize()
    # replay again: same result (workspace is left clean, buffers are reused)
    assert np.array_equal(host(st.decode_step()), logits)
    assert st.launches_per_step == (2 * 4 + 1 if fuse else 2 * 7 + 1)

    x = x0
    for li in range(2):
        w = st.layers[li]
        q = _oracle_linear(w["q"], x, fmt)
        x1 = _oracle_linear(w["o"], q, fmt, qo.EP_ADD, x)
        g = _oracle_linear(w["gate"], x1, fmt, qo.EP_SILU)
        a = _oracle_linear(w["up"], x1, fmt, qo.EP_MULTIPLY, g)
        x = _oracle_linear(w["down"], a, fmt, qo.EP_ADD, x1)
    ref = _oracle_linear(st.lm_head, x, fmt, qo.EP_CAST_F32)
    assert np.isfinite(ref).all()
    scale = np.abs(ref).max()
    assert np.abs(logits - ref).max() <= 2e-2 * scale, (np.abs(logits - ref).max(), scale)


def test_neox_stack_f32_matches_oracle_chain():
    name = "q4f32_0"
    fmt = qo.FORMATS[name]
    shape = ModelShape("tiny-neox", "gpt_neox", 512, 2048, 2, 1000)
    st = DecodeStack(shape, name, layers=2)
    x0 = (np.random.default_rng(1).standard_normal(shape.hidden) * 0.5).astype(np.float32)
    st.x_in.copy_(torch.from_numpy(x0))
    st.capture()
    logits = host(st.decode_step().clone())
    x = x0
    for li in range(2):
        w, b = st.layers[li], {k: host(v) for k, v in st.biases[li].items()}
        q = _oracle_linear(w["q"], x, fmt, qo.EP_BIAS, b["q"])
        x1 = _oracle_linear(w["dense"], q, fmt, qo.EP_BIAS, b["dense"])
        a = _oracle_linear(w["h_to_4h"], x1, fmt, qo.EP_BIAS_GELU, b["h_to_4h"])
        x = _oracle_linear(w["4h_to_h"], a, fmt, qo.EP_BIAS, b["4h_to_h"])
    ref = _oracle_linear(st.lm_head, x, fmt, qo.EP_NONE)
    assert np.abs(logits - ref).max() <= 1e-4 * max(np.abs(ref).max(), 1e-3)


def test_host_entry_round_trip():
    shape = ModelShape("tiny-llama", "llama", 512, 1408, 2, 1000)
    st = DecodeStack(shape, "q4f16_0", layers=2)
    st.capture()
    xh = (torch.randn(shape.hidden) * 0.5).to(torch.float16).pin_memory()
    lh = torch.empty(shape.vocab, dtype=torch.float32).pin_memory()
    st.decode_step_host(xh, lh)
    st.x_in.copy_(xh)
    assert torch.equal(st.decode_step().cpu(), lh)
