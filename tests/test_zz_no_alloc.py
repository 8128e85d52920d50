"""b200q_gemv_static (-m gpu): the GEMV entry point that allocates nothing - its split-K scratch is a static device
array inside the library (include/b200q.h).  It has no counterpart in the reference;
the checker here is the existing entry point b200q_gemv on the same inputs, plus the oracle's tolerance.

Collected last on purpose (the name sorts behind every other test file).
The test passes alone in a fresh process and after the other tests (the array keeps whatever the last call left).
"""
import ctypes

import numpy as np
import pytest
import torch

from mlc_llm_b200 import _lib, ops
from oracle import quant_oracle as qo
from util import beq, check_gemv_tolerance, dev, host, rand_weight

pytestmark = pytest.mark.gpu


def _inputs(name, K, N, seed):
    fmt = qo.FORMATS[name]
    packed, scale = rand_weight(fmt, K, N, seed=seed)
    rng = np.random.default_rng(seed + 1)
    x = (rng.standard_normal(K) * 0.5).astype(fmt.np_dtype)
    return fmt, packed, scale, x


def _call_static(p, s, x, out, K, N, fmt_id, ep=0):
    """straight through the C-ABI: pointers and sizes only, outputs allocated by the caller beforehand"""
    L = _lib.lib()
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    return L.b200q_gemv_static(p.data_ptr(), s.data_ptr(), x.data_ptr(), None, out.data_ptr(), K, N, fmt_id, ep, 0, stream)


# two calls in a row, different shapes and inputs: the second one finds the first one's partial sums in the array
@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0", "q4f32_0", "q8f16_0"])
def test_static_scratch_matches_workspace_entry_point(name):
    spec = ops._spec(name)
    for call, (K, N) in enumerate([(4096, 4096), (11008, 512)]):  # N = 512: every panel is split over many CTAs
        fmt, packed, scale, x = _inputs(name, K, N, seed=100 * call + K + N)
        p, s, xd = dev(packed), dev(scale), dev(x)
        want = ops.gemv(p, s, xd, name)  # b200q_gemv with a zero-filled caller-provided workspace
        out = torch.full_like(want, float("nan"))
        before = torch.cuda.memory_allocated()
        rc = _call_static(p, s, xd, out, K, N, spec.fmt_id)
        assert rc == 0, _lib.lib().b200q_last_error().decode()
        assert torch.cuda.memory_allocated() == before
        torch.cuda.synchronize()
        got = host(out)
        # same kernel, same split, partial sums added in CTA order: the two entry points agree to the bit; the bar
        # the other GEMV tests use is checked as well
        assert beq(got, host(want)), (name, K, N, call)
        check_gemv_tolerance(got, packed, scale, x, fmt, out_f32=(want.dtype == torch.float32), what=f"static {name} {K}x{N}")


def test_static_scratch_twice_same_inputs_and_epilogue():
    name, K, N = "q4f16_0", 4096, 1024
    fmt, packed, scale, x = _inputs(name, K, N, seed=7)
    p, s, xd = dev(packed), dev(scale), dev(x)
    ep0 = (torch.randn(N, device="cuda") * 0.1).half()
    want = ops.gemv(p, s, xd, name, epilogue="add", ep0=ep0)
    a = ops.gemv_static(p, s, xd, name, epilogue="add", ep0=ep0)
    b = ops.gemv_static(p, s, xd, name, epilogue="add", ep0=ep0)
    torch.cuda.synchronize()
    assert beq(host(a), host(want)) and beq(host(b), host(want))


def test_static_scratch_under_graph_capture():
    """the use case: captured into a CUDA graph (no allocation is legal there), replayed twice"""
    name, K, N = "q4f16_0", 4096, 512
    spec = ops._spec(name)
    fmt, packed, scale, x = _inputs(name, K, N, seed=11)
    p, s, xd = dev(packed), dev(scale), dev(x)
    want = ops.gemv(p, s, xd, name)
    out = torch.zeros_like(want)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        rc = _call_static(p, s, xd, out, K, N, spec.fmt_id)
    assert rc == 0, _lib.lib().b200q_last_error().decode()
    for _ in range(2):
        out.zero_()
        g.replay()
        torch.cuda.synchronize()
        assert beq(host(out), host(want))


def test_static_scratch_rejects_beyond_its_maximum():
    L = _lib.lib()
    spec = ops._spec("q4f16_0")
    # (never dereferenced: the call is refused before any launch)
    w = torch.zeros(64, dtype=torch.int32, device="cuda")
    N = 4096 * 128 + 128
    sc, x = (torch.zeros(64, dtype=torch.float16, device="cuda") for _ in range(2))
    out = torch.zeros(N, dtype=torch.float16, device="cuda")  # (full size: the alias check looks at out's whole range)
    rc = L.b200q_gemv_static(w.data_ptr(), sc.data_ptr(), x.data_ptr(), None, out.data_ptr(), 32, N, spec.fmt_id, 0, 0, None)
    assert rc == -1, rc  # B200Q_ERR_BAD_SHAPE
    msg = L.b200q_last_error().decode()
    assert "static scratch" in msg and "4096" in msg, msg
    # formats without a split-K kernel have no scratch to replace
    v = ops._spec("q3f16_u16")
    rc = L.b200q_gemv_static(w.data_ptr(), sc.data_ptr(), x.data_ptr(), None, out.data_ptr(), 40, 8, v.fmt_id, 0, 0, None)
    assert rc != 0
