"""GPU tests of the persistent decode-program kernel (b200q_gemv_program_*): same results as the
step-by-step launches, the oracle tolerance, dependency handling, replay."""
import numpy as np
import pytest
import torch

from mlc_llm_b200 import ops
from mlc_llm_b200.decode_stack import DecodeStack
from oracle import quant_oracle as qo
from util import beq, check_gemv_tolerance, dev, host, rand_weight

pytestmark = pytest.mark.gpu


def _w(fmt, K, N, seed):
    p, s = rand_weight(fmt, K, N, seed)
    s = (s.astype(np.float32) * (3.0 / ((1 << (fmt.nbit - 1)) - 1))).astype(s.dtype)  # same weight range for every bit width
    return p, s, dev(p), dev(s)


@pytest.mark.parametrize("name", list(qo.FORMATS))
def test_single_step_matches_gemv_and_oracle(name):
    fmt = qo.FORMATS[name]
    f32 = fmt.dtype == "float32"
    for K, N in ((4096, 4096), (1376, 520), (96 if fmt.nbit != 3 else 120, 8), (11008, 4096)):
        p, s, pd, sd = _w(fmt, K, N, seed=K + N)
        x = qo.synth_act(K, fmt, seed=3)
        xd = dev(x)
        out = torch.zeros(N, dtype=torch.float32 if f32 else torch.float16, device="cuda")
        prog = ops.GemvProgram([dict(weights=[(pd, sd)], x=xd, outs=[out])], name)
        prog.run()
        y = host(out)
        tc = ops.program_uses_tensor_cores(name)  # tcgen05 program: group scale applied to the f32 group sum
        check_gemv_tolerance(y, p, s, x, fmt, out_f32=f32, what=f"{name} {K}x{N}", group_scaled=tc)
        y1 = host(ops.gemv(pd, sd, xd, name)).reshape(-1)
        if tc:
            # same value up to the per-weight f16 rounding the single-op kernel reproduces: a few f16 ulps
            d = np.abs(y.astype(np.float64) - y1.astype(np.float64))
            assert d.max() <= 2.0**-8 * np.abs(y1.astype(np.float64)).max()
        elif not f32:
            assert (np.abs(y.view(np.int16).astype(np.int32) - y1.view(np.int16).astype(np.int32)) <= 1).mean() > 0.99
        out.zero_()
        prog.run()  # replay: the workspace must have been left clean
        assert beq(host(out), y)


def test_unaligned_plain_input_q4f16_0():
    """A plain (non-LL) activation that is not 16-byte aligned, with a K tail: the element-wise path of the
    activation warps of the tcgen05 program."""
    fmt = qo.FORMATS["q4f16_0"]
    for K, N in ((1000, 264), (4096, 128)):
        p, s, pd, sd = _w(fmt, K, N, seed=K)
        x = qo.synth_act(K, fmt, seed=5)
        buf = torch.zeros(K + 8, dtype=torch.float16, device="cuda")
        xd = buf[3:3 + K]
        xd.copy_(dev(x))
        assert xd.data_ptr() % 16 != 0
        out = torch.zeros(N, dtype=torch.float16, device="cuda")
        ops.GemvProgram([dict(weights=[(pd, sd)], x=xd, outs=[out])], "q4f16_0").run()
        check_gemv_tolerance(host(out), p, s, x, fmt, out_f32=False, what=f"unaligned {K}x{N}", group_scaled=True)


@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0", "q8f16_0", "q4f32_0"])
def test_chain_equals_step_by_step(name):
    """A Llama layer's chain (q|k|v -> o+res -> gate|up -> down(x*x_mul)+res) twice: program == launches."""
    fmt = qo.FORMATS[name]
    dt = torch.float32 if fmt.dtype == "float32" else torch.float16
    h, inter = 512, 1376 if fmt.nbit != 3 else 1360
    z = lambda n: torch.zeros(n, dtype=dt, device="cuda")
    x0 = dev(qo.synth_act(h, fmt, seed=1, std=0.5))
    steps, ref_launches = [], []
    x = x0
    for layer in range(2):
        W = {k: _w(fmt, K, N, seed=layer * 10 + i)[2:] for i, (k, K, N) in enumerate(
            [("q", h, h), ("k", h, h), ("v", h, h), ("o", h, h), ("gate", h, inter), ("up", h, inter), ("down", inter, h)])}
        b = {k: z(n) for k, n in [("q", h), ("k", h), ("v", h), ("x1", h), ("g", inter), ("u", inter), ("x2", h)]}
        r = {k: z(n) for k, n in [("q", h), ("k", h), ("v", h), ("x1", h), ("g", inter), ("u", inter), ("x2", h)]}
        steps += [
            dict(weights=[W["q"], W["k"], W["v"]], x=x, outs=[b["q"], b["k"], b["v"]]),
            dict(weights=[W["o"]], x=b["q"], epilogues=["add"], ep0s=[x], outs=[b["x1"]]),
            dict(weights=[W["gate"], W["up"]], x=b["x1"], epilogues=["silu", "none"], outs=[b["g"], b["u"]]),
            dict(weights=[W["down"]], x=b["g"], x_mul=b["u"], epilogues=["add"], ep0s=[b["x1"]], outs=[b["x2"]]),
        ]
        ref_launches.append((W, r))
        x = b["x2"]
    prog = ops.GemvProgram(steps, name)
    prog.run()
    torch.cuda.synchronize()
    got = x.clone()
    # step by step with the single-op kernels
    xr = x0
    for W, r in ref_launches:
        ops.gemv_multi([W["q"], W["k"], W["v"]], xr, name, outs=[r["q"], r["k"], r["v"]])
        ops.gemv(*W["o"], r["q"], name, "add", xr, out=r["x1"])
        ops.gemv_multi([W["gate"], W["up"]], r["x1"], name, epilogues=["silu", "none"], outs=[r["g"], r["u"]])
        ops.gemv_multi([W["down"]], r["g"], name, epilogues=["add"], ep0s=[r["x1"]], outs=[r["x2"]], x_mul=r["u"])
        xr = r["x2"]
    a, e = host(got).astype(np.float64), host(xr).astype(np.float64)
    assert np.isfinite(a).all()
    # different split-K grouping of the same exact products: equal to f16 rounding noise through two layers
    assert np.abs(a - e).max() <= 2e-2 * max(1.0, np.abs(e).max())
    for _ in range(3):  # replays are bit-identical (deterministic reduction order)
        prog.run()
        torch.cuda.synchronize()
        assert beq(host(x), host(got))


def test_program_in_cuda_graph_and_full_stack():
    stack = DecodeStack("llama-7b", "q4f16_0", layers=2, use_program=True)
    stack.x_in.copy_((torch.randn(4096, device="cuda") * 0.05).half())
    stack.capture()
    y1 = stack.decode_step().clone()
    y2 = stack.decode_step().clone()
    assert torch.equal(y1, y2) and torch.isfinite(y1).all()
    ref = DecodeStack("llama-7b", "q4f16_0", layers=2, use_program=False)
    ref.x_in.copy_(stack.x_in)
    ref.capture()
    y3 = ref.decode_step()
    assert (y1 - y3).abs().max().item() <= 2e-2 * max(1.0, y3.abs().max().item())


def test_launch_tag_wraps_around():
    """The LL tag is (launch epoch % 65535) + 1: more than 65535 replays of one program must keep giving the
    same bits (every word is rewritten by every launch, so a tag can never be confused with a stale one)."""
    stack = DecodeStack("llama-7b", "q4f16_0", layers=1, use_program=True)
    stack.x_in.copy_((torch.randn(4096, device="cuda") * 0.05).half())
    stack.capture()
    y0 = stack.decode_step().clone()
    for i in range(66000):
        stack.decode_step()
    torch.cuda.synchronize()
    assert torch.equal(stack.decode_step(), y0)
