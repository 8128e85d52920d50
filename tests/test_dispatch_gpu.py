"""The dispatch seam (mlc_llm_b200/dispatch.py), every branch: the contract of the reference's lookup_func
(mlc_llm/dispatch/llama/main.py:6452-6458; dispatch_tir_operator.py:22-33) - an exact (scheme, K, N, epilogue)
key returns a destination-passing operator f(packed, scale, x[, operands], out); a miss returns None."""
import numpy as np
import pytest
import torch

from mlc_llm_b200 import dispatch, ops
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from util import beq, check_gemv_tolerance, dev, host, rand_weight

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(qo.FORMATS))
def test_lookup_gemv_branch_every_epilogue(name):
    fmt = qo.FORMATS[name]
    f32 = fmt.dtype == "float32"
    K, N = (2560, 512) if fmt.nbit != 3 else (2560, 512)
    packed, scale = rand_weight(fmt, K, N, seed=31)
    x, epv = qo.synth_act(K, fmt, seed=32), qo.synth_act(N, fmt, seed=33)
    p, s, xd, ed = dev(packed), dev(scale), dev(x), dev(epv)
    dt = torch.float32 if f32 else torch.float16
    y_none = None
    for ep in dispatch.GEMV_EPILOGUES:
        op = dispatch.lookup(name, K, N, ep, m_is_one=True)
        assert op is not None and op.__name__.startswith("fused_decode_")
        out = torch.zeros(N, dtype=torch.float32 if (ep == "cast_f32" or f32) else dt, device="cuda")
        operands = [ed] if dispatch.N_OPERANDS[ep] else []
        op(p, s, xd, *operands, out)  # PrimFunc argument order: (packed, scale, x[, operand], out)
        y = host(out)
        if ep == "none":
            y_none = y
            check_gemv_tolerance(y, packed, scale, x, fmt, out_f32=f32, what=f"lookup {name}")
        elif ep in ("add", "multiply", "bias") and not f32:
            assert beq(y, qo.apply_epilogue(y_none, qo.EPILOGUES[ep], epv)), ep
        with pytest.raises(TypeError):
            op(p, s, xd, ed, ed, out)  # wrong operand count
    # misses: unknown scheme / epilogue, N not a multiple of 8, a GEMM-only epilogue on the GEMV branch
    assert dispatch.lookup("q5f16_9", K, N) is None
    assert dispatch.lookup(name, K, N, "tanh") is None
    assert dispatch.lookup(name, K, 12) is None
    assert dispatch.lookup(name, K, N, "bias_add2", m_is_one=True) is None


def test_lookup_gemm_branch_and_f32_miss():
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    M, K, N = 64, 512, 256
    packed, scale = rand_weight(fmt, K, N, seed=41)
    X = qo.synth_act((M, K), fmt, seed=42)
    op = dispatch.lookup(name, K, N, "none", m_is_one=False)
    assert op is not None and "NT_matmul" in op.__name__
    Y = torch.zeros(M, N, dtype=torch.float16, device="cuda")
    op(dev(packed), dev(scale), dev(X), Y)
    y64, ab = qo.gemm_f64(packed, scale, X, fmt)
    assert (np.abs(host(Y).astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all()
    assert dispatch.lookup("q4f32_0", K, N, "none", m_is_one=False) is None  # f32 prefill: the reference's miss


@pytest.mark.parametrize("name", list(qo.FORMATS))
def test_lookup_decode_both_layouts(name):
    fmt = qo.FORMATS[name]
    K, N = (1280, 264) if fmt.nbit != 3 else (1280, 264)
    packed, scale = rand_weight(fmt, K, N, seed=51)
    for tr in (False, True):
        op = dispatch.lookup_decode(name, K, N, tr)
        assert op is not None and op.__name__.startswith("decode_")
        out = torch.zeros((N, K) if tr else (K, N), dtype=torch.float32 if fmt.dtype == "float32" else torch.float16, device="cuda")
        op(dev(packed), dev(scale), out)
        assert beq(host(out), co.decode(packed, scale, K, fmt, transpose_output=tr))
    assert dispatch.lookup_decode("nope", K, N, False) is None
    assert dispatch.lookup_decode(name, K, 12, False) is None


def test_lookup_program_returns_the_decode_program():
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    h, inter = 512, 1408
    z = lambda n, d=torch.float16: torch.zeros(n, dtype=d, device="cuda")
    W = {k: tuple(dev(a) for a in rand_weight(fmt, K, N, seed=i)) for i, (k, K, N) in enumerate(
        [("q", h, h), ("k", h, h), ("v", h, h), ("o", h, h), ("gate", h, inter), ("up", h, inter), ("down", inter, h)])}
    x0 = dev(qo.synth_act(h, fmt, seed=1, std=0.5))
    b = {k: z(n) for k, n in [("q", h), ("k", h), ("v", h), ("x1", h), ("g", inter), ("u", inter), ("x2", h)]}
    sites = [
        dict(weights=[W["q"], W["k"], W["v"]], x=x0, outs=[b["q"], b["k"], b["v"]]),
        dict(weights=[W["o"]], x=b["q"], epilogues=["add"], ep0s=[x0], outs=[b["x1"]]),
        dict(weights=[W["gate"], W["up"]], x=b["x1"], epilogues=["silu", "none"], outs=[b["g"], b["u"]]),
        dict(weights=[W["down"]], x=b["g"], x_mul=b["u"], epilogues=["add"], ep0s=[b["x1"]], outs=[b["x2"]]),
    ]
    prog = dispatch.lookup_program(name, sites)
    assert isinstance(prog, ops.GemvProgram) and prog.n_steps == 4
    prog.run()
    torch.cuda.synchronize()
    assert torch.isfinite(b["x2"]).all() and b["x2"].abs().max().item() > 0
    bad = [dict(weights=[W["q"]], x=x0, epilogues=["bias_add2"], outs=[b["q"]])]
    assert dispatch.lookup_program(name, bad) is None
    assert dispatch.lookup_program("nope", sites) is None
