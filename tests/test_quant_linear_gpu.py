"""The caller-side modules (mlc_llm_b200/quant_linear.py): the GroupQuantize rewrite of a torch module tree
(quantization.py:362-420: which parameters are packed, decode-step vs prefill operator per call site, take_decode
for the embedding) and loading a reference-format artefact directory (utils.py:113-139) into them."""
import numpy as np
import pytest
import torch

from mlc_llm_b200 import param_cache
from mlc_llm_b200.quant_linear import QuantEmbedding, QuantLinear, from_param_cache, quantize_model
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from util import beq, host

pytestmark = pytest.mark.gpu


class TinyMLP(torch.nn.Module):
    def __init__(self, h=256, inter=512, vocab=300):
        super().__init__()
        self.embed = torch.nn.Embedding(vocab, h)
        self.gate = torch.nn.Linear(h, inter, bias=False)
        self.up = torch.nn.Linear(h, inter, bias=False)
        self.down = torch.nn.Linear(inter, h, bias=True)
        self.norm = torch.nn.LayerNorm(h)


def _decode(fmt, mod):
    p, s = host(mod.packed), host(mod.scale)
    if fmt.sym:
        return co.decode(p, s, mod.K, fmt)
    return co.decode_asym(p, s, None if mod.zmin is None else host(mod.zmin), mod.K, fmt)


@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0", "q4f16_asym", "q4f32_0"])
def test_quantize_model_rewrite_and_call_sites(name):
    fmt = qo.ALL_FORMATS[name]
    torch.manual_seed(0)
    h, inter = (256, 512) if fmt.nbit != 3 else (320, 640)
    m = TinyMLP(h, inter).cuda().to(torch.float16 if fmt.dtype == "float16" else torch.float32)
    W_gate = m.gate.weight.detach().clone()
    quantize_model(m, name)
    assert isinstance(m.gate, QuantLinear) and isinstance(m.down, QuantLinear) and isinstance(m.embed, QuantEmbedding)
    assert isinstance(m.norm, torch.nn.LayerNorm)  # not a matmul / take parameter: left alone
    # the packed tensors are the reference encoder's, bit for bit
    enc = qo.encode_sym(host(W_gate), fmt) if fmt.sym else qo.encode_asym(host(W_gate), fmt)
    assert beq(host(m.gate.packed), enc[0]) and beq(host(m.gate.scale), enc[1])
    wd = _decode(fmt, m.gate).astype(np.float64)
    dt = m.gate.scale.dtype if fmt.sym or fmt.dtype == "float16" else torch.float32
    f32 = fmt.dtype == "float32"
    # decode step (x.shape[-2] == 1): fused dequant-GEMV
    x1 = (torch.randn(1, 1, h, device="cuda") * 0.5).to(torch.float32 if f32 else torch.float16)
    y1 = host(m.gate(x1)).reshape(-1).astype(np.float64)
    t1 = host(x1).reshape(-1).astype(np.float64) @ wd
    ab = np.abs(host(x1).reshape(-1).astype(np.float64)) @ np.abs(wd)
    assert (np.abs(y1 - t1) <= qo.tolerance(t1, ab, f32)).all()
    # prefill (M > 1): dequant-GEMM (f16) / decode + matmul (f32, the dispatch miss)
    xm = (torch.randn(1, 130, h, device="cuda") * 0.5).to(x1.dtype)
    ym = host(m.gate(xm))[0].astype(np.float64)
    tm = host(xm)[0].astype(np.float64) @ wd
    abm = np.abs(host(xm)[0].astype(np.float64)) @ np.abs(wd)
    tol = qo.tolerance(tm, abm, f32) * (4 if f32 else 1)  # (torch's f32 matmul may use a different summation order)
    assert (np.abs(ym - tm) <= tol).all()
    # bias epilogue by default when the layer has one; fused silu on request
    a = (torch.randn(1, 1, inter, device="cuda") * 0.5).to(x1.dtype)
    yb = host(m.down(a)).reshape(-1)
    y0 = host(m.down(a, epilogue="none")).reshape(-1)
    assert beq(yb, qo.apply_epilogue(y0, qo.EP_BIAS, host(m.down.bias)))
    # embedding: take_decode
    ids = torch.tensor([[3, 0, 299, 3]], device="cuda")
    e = host(m.embed(ids))
    pr, sr = host(m.embed.packed_rm), host(m.embed.scale_rm)
    if fmt.sym:
        ref = qo.take_decode_sym(pr, sr, host(ids).reshape(-1), h, fmt)
    else:
        ref = qo.take_decode_asym(pr, sr, None if m.embed.zmin_rm is None else host(m.embed.zmin_rm), host(ids).reshape(-1), h, fmt)
    assert beq(e.reshape(-1, h), ref)


def test_from_param_cache_loads_a_reference_artefact(tmp_path):
    """dump (= save_params) a quantized parameter list, load it the way load_params does, run the modules."""
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    emb = qo.synth_weight(120, 256, fmt, seed=1)
    W = qo.synth_weight(512, 256, fmt, seed=2)
    pe, se = qo.encode_sym_rowmajor(emb, fmt)
    pw, sw = qo.encode_sym(W, fmt)
    norm = np.ones(256, np.float16)
    d = str(tmp_path / "params")
    param_cache.dump([pe, se, pw, sw, norm], d)
    mods = from_param_cache(d, name)
    assert isinstance(mods[0], QuantEmbedding) and isinstance(mods[1], QuantLinear) and isinstance(mods[2], torch.Tensor)
    x = mods[0](torch.tensor([5], device="cuda"))           # [1, 256]
    assert beq(host(x), qo.take_decode_sym(pe, se, np.array([5]), 256, fmt))
    y = host(mods[1](x.view(1, 1, 256))).reshape(-1)
    y64, ab = co.gemv_f64(pw, sw, host(x).reshape(-1), fmt)
    assert (np.abs(y.astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all()
