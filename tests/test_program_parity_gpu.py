"""Parity of the decode-PROGRAM kernels (b200q_gemv_program_*; for q4f16_0 the tcgen05 kernel that carries
bench.py's headline) against the reference-run goldens and the oracle: the same cases the single-op stream
kernel gets in test_gpu_parity.py, plus the numerics of the tcgen05 formulation itself.

The tcgen05 program evaluates  y = sum_g s[g,n] * (T[g,n] - C[g]),  T = sum_k (off_k + q) x  on the tensor cores,
C = sum_k (off_k + 7) x: the group scale multiplies an f32 group sum instead of every weight being rounded to
f16 first.  Claims tested here:
  * against the f64 sum of the UNROUNDED (q - 7) * s it meets the ORIGINAL SURVEY 8c(ii) bound
    (2^-10 |y| + 2^-14 sum|x w| + 2^-24) for zero-mean activations at every Llama shape - "closer to real
    arithmetic than the reference's f16((q-7)*s) weights";
  * the  T - C  cancellation is bounded by 2^-22 * sum_k off_k |x_k| s (quant_oracle.tolerance_group_scaled) on
    adversarial inputs: all-q=7 columns, activations with a large mean, outlier channels, |x| near the f16 maximum.
"""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from mlc_llm_b200 import ops
from mlc_llm_b200.decode_stack import DecodeStack, ModelShape
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from oracle import stack_check
from oracle.golden_inputs import epilogue_of, sha, synth_fused_inputs
from util import beq, dev, host, rand_weight

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EP_NAME = {v: k for k, v in qo.EPILOGUES.items()}


def run_program(name, packed, scale, x, ep="none", epv=None):
    """One GEMV as a one-step program (the program kernel's own staging, reduction and epilogue code)."""
    fmt = qo.FORMATS[name]
    N = packed.shape[1]
    f32 = fmt.dtype == "float32" or ep == "cast_f32"
    out = torch.zeros(N, dtype=torch.float32 if f32 else torch.float16, device="cuda")
    step = dict(weights=[(dev(packed), dev(scale))], x=dev(x), epilogues=[ep],
                ep0s=[dev(epv) if epv is not None else None], outs=[out])
    ops.GemvProgram([step], name).run()
    torch.cuda.synchronize()
    return host(out)


def matmul_bound(name, packed, scale, x, out_f32, K=None):
    """(truth, tolerance) of the matmul for the program kernel of this format."""
    fmt = qo.FORMATS[name]
    if ops.program_uses_tensor_cores(name):
        y64, ab, cc = co.gemv_f64_unrounded(packed, scale, x, fmt, K)
        return y64, qo.tolerance_group_scaled(y64, ab, cc, out_f32), qo.tolerance(y64, ab, out_f32)
    y64, ab = co.gemv_f64(packed, scale, x, fmt, K)
    t = qo.tolerance(y64, ab, out_f32)
    return y64, t, t


@pytest.fixture(scope="module")
def fused(golden_dir):
    return np.load(os.path.join(golden_dir, "fused_gemv.npz"))


def test_program_against_reference_goldens(fused):
    """The 12 reference-run fused PrimFunc outputs (dispatch_tir_operator_adreno.py:5-411 q4f16_0,
    dispatch/llama/main.py:5316-5758 int3) through the PROGRAM kernels: inside the stated tolerance of the f64
    truth and closer to it than the reference's own f16-accumulated result."""
    names = sorted({k.rsplit("/", 1)[0] for k in fused.files})
    assert len(names) == 12
    for n in names:
        fname = n.split("/")[0]
        fmt = qo.FORMATS[fname]
        K, N, seed, _ = [int(v) for v in fused[n + "/meta"]]
        packed, scale, x, epv = synth_fused_inputs(fmt, K, N, seed)
        assert sha(packed, scale, x, epv) == bytes(fused[n + "/sha"]).decode()
        ep = epilogue_of(n)
        need = ep in (qo.EP_ADD, qo.EP_MULTIPLY)
        y = run_program(fname, packed, scale, x, EP_NAME[ep], epv if need else None).reshape(-1)
        y_ref = fused[n + "/y"]
        assert y.dtype == y_ref.dtype and y.shape == y_ref.shape
        truth, tol = stack_check.op_truth_and_tol(packed, scale, x, fmt, ep, epv if need else None, K, ops.program_uses_tensor_cores(fname))
        err = np.abs(y.astype(np.float64) - truth)
        assert (err <= tol).all(), (n, float((err / tol).max()))
        # closer to the truth of the reference's own (rounded) weights than the reference is
        y64, _ = co.gemv_f64(packed, scale, x, fmt)
        t64 = {qo.EP_ADD: lambda: epv.astype(np.float64) + y64, qo.EP_MULTIPLY: lambda: epv.astype(np.float64) * y64,
               qo.EP_SILU: lambda: y64 / (1.0 + np.exp(-y64))}.get(ep, lambda: y64)()
        ours, ref = np.abs(y.astype(np.float64) - t64).max(), np.abs(y_ref.astype(np.float64) - t64).max()
        assert ours <= ref, (n, ours, ref)


SHAPES = [(4096, 4096), (4096, 11008), (11008, 4096), (4096, 32000), (2560, 2560), (10240, 2560), (5120, 13824),
          (1376, 4096), (8192, 1024), (64, 8), (1000, 264), (96, 264)]


@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0", "q8f16_0", "q4f32_0"])
@pytest.mark.parametrize("K,N", SHAPES)
def test_program_tolerance_shapes(name, K, N):
    """Every Llama / NeoX GEMV shape (incl. 4096->11008, 4096->32000 and K tails) through the program kernels,
    N(0,1) activations.  q4f16_0 (tcgen05): the ORIGINAL bound against the unrounded truth."""
    fmt = qo.FORMATS[name]
    if fmt.nbit == 3:
        K = -(-K // 40) * 40 if K > 100 else 120
    packed, scale = rand_weight(fmt, K, N, seed=5 * K + N)
    x = qo.synth_act(K, fmt, seed=K + 1)
    f32 = fmt.dtype == "float32"
    y = run_program(name, packed, scale, x).reshape(-1)
    y64, tol, tol_orig = matmul_bound(name, packed, scale, x, f32)
    err = np.abs(y.astype(np.float64) - y64)
    assert (err <= tol).all(), float((err / tol).max())
    assert (err <= tol_orig).all(), ("original 2^-14 bound", float((err / tol_orig).max()))


def test_program_lm_head_cast_f32_and_truth_ordering():
    """4096 -> 32000 with the f32 (lm_head) epilogue on the tcgen05 program, encoder-produced weights: f32-output
    tolerance against the unrounded truth, and the claim itself - the program is at least as close to the
    unrounded real-arithmetic value as the exact per-weight kernel (which reproduces the reference's f16 weights)."""
    name, K, N = "q4f16_0", 4096, 32000
    fmt = qo.FORMATS[name]
    W = qo.synth_weight(N, K, fmt, seed=77)
    packed, scale = qo.encode_sym(W, fmt)
    x = qo.synth_act(K, fmt, seed=78)
    y = run_program(name, packed, scale, x, "cast_f32").reshape(-1)
    assert y.dtype == np.float32
    y64u, ab, cc = co.gemv_f64_unrounded(packed, scale, x, fmt)
    err = np.abs(y.astype(np.float64) - y64u)
    assert (err <= qo.tolerance_group_scaled(y64u, ab, cc, True)).all()
    y_stream = host(ops.gemv(dev(packed), dev(scale), dev(x), name, "cast_f32")).reshape(-1)
    err_stream = np.abs(y_stream.astype(np.float64) - y64u)
    assert np.sqrt((err**2).mean()) <= np.sqrt((err_stream**2).mean()), (err.max(), err_stream.max())


@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0", "q8f16_0"])
def test_program_epilogues_op_for_op(name):
    """Every epilogue of the program kernels equals the oracle's dtype-typed epilogue applied to the program's own
    un-fused result (bit-exact for add / multiply / bias, <= 1 f16 ulp where exp / erf are involved)."""
    fmt = qo.FORMATS[name]
    K, N = (4096, 1024) if fmt.nbit != 3 else (4000, 1024)
    packed, scale = rand_weight(fmt, K, N, seed=9)
    x = qo.synth_act(K, fmt, seed=10)
    epv = qo.synth_act(N, fmt, seed=12)
    y0 = run_program(name, packed, scale, x).reshape(-1)
    for ep in ("add", "multiply", "bias"):
        y = run_program(name, packed, scale, x, ep, epv).reshape(-1)
        assert beq(y, qo.apply_epilogue(y0, qo.EPILOGUES[ep], epv)), ep
    for ep in ("silu", "bias_gelu"):
        y = run_program(name, packed, scale, x, ep, epv if ep == "bias_gelu" else None).reshape(-1)
        e = qo.apply_epilogue(y0, qo.EPILOGUES[ep], epv)
        ulp = np.abs(y.view(np.int16).astype(np.int32) - e.view(np.int16).astype(np.int32))
        assert ulp.max() <= 1, (ep, ulp.max())
    y32 = run_program(name, packed, scale, x, "cast_f32").reshape(-1)
    assert y32.dtype == np.float32 and beq(y32.astype(np.float16), y0)


def _adversarial_cases(K, N, rng):
    fmt = qo.FORMATS["q4f16_0"]
    packed, scale = rand_weight(fmt, K, N, seed=123)
    base = rng.standard_normal(K).astype(np.float32)
    cases = {}
    # (a) columns whose codes are all 7 (weight exactly zero): the reference gives exactly 0
    p7 = packed.copy()
    p7[:, : N // 2] = np.uint32(0x77777777)
    cases["all_q7_columns"] = (p7, scale, base.astype(np.float16))
    # (b) large mean: x = 80 + N(0,1)  (post-residual streams of real models)
    cases["mean_80"] = (packed, scale, (80.0 + base).astype(np.float16))
    # (c) four outlier channels of +-200
    xo = base.copy()
    xo[[3, 1001 % K, K // 2, K - 5]] = [200.0, -200.0, 200.0, -200.0]
    cases["outliers_200"] = (packed, scale, xo.astype(np.float16))
    # (d) all-positive activations (post-SiLU * up products)
    cases["all_positive"] = (packed, scale, np.abs(base * 3.0).astype(np.float16))
    # (e) |x| at 0.9 * f16 max with tiny scales (the products (1024+q) x reach 6e7, the sums 1e9)
    big = (np.sign(base) * 0.9 * 65504.0).astype(np.float16)
    tiny = (scale.astype(np.float32) * 2.0**-10).astype(np.float16)
    cases["near_f16_max"] = (packed, tiny, big)
    return cases


@pytest.mark.parametrize("K,N", [(4096, 512), (11008, 256)])
def test_tc_program_offset_cancellation_adversarial(K, N):
    """The 1024 + q / 64 + q offset trick cancels two ~10^3 * sum|x| terms in f32: bounded by
    2^-22 * sum_k off_k |x_k| s on top of the standard terms, on inputs built to make it hurt."""
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    rng = np.random.default_rng(K)
    report = {}
    for label, (packed, scale, x) in _adversarial_cases(K, N, rng).items():
        y = run_program(name, packed, scale, x, "cast_f32").reshape(-1)
        assert np.isfinite(y).all(), label
        y64u, ab, cc = co.gemv_f64_unrounded(packed, scale, x, fmt)
        err = np.abs(y.astype(np.float64) - y64u)
        tol = qo.tolerance_group_scaled(y64u, ab, cc, True)
        report[label] = float((err / tol).max())
        assert (err <= tol).all(), (label, report[label], float(err.max()))
        if label == "all_q7_columns":
            # no contribution of the weights survives: only the cancellation residue (the exact kernel returns 0)
            z = np.abs(y[: N // 2].astype(np.float64))
            assert (z <= 2.0**-22 * cc[: N // 2] + 2.0**-24).all(), float(z.max())
            y_exact = host(ops.gemv(dev(packed), dev(scale), dev(x), name, "cast_f32")).reshape(-1)
            assert (y_exact[: N // 2] == 0).all()
        # RMS error against the unrounded truth: never worse than a few 2^-11 of the output RMS
        rms = np.sqrt((err**2).mean()) / max(np.sqrt((y64u**2).mean()), 1e-30)
        assert rms <= 4 * 2.0**-11, (label, rms)
    print("tc program adversarial err/tol:", report)


def test_exact_program_path_documented_env():
    """B200Q_PROG_TC=0 selects the exact per-weight program kernel for q4f16_0 (the documented bit-faithful
    path): its result meets the standard bound against the ROUNDED-weight truth and all-q=7 columns give 0."""
    code = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r + "/tests")
from mlc_llm_b200 import ops
from oracle import c_oracle as co, quant_oracle as qo
from util import dev, host, rand_weight
fmt = qo.FORMATS["q4f16_0"]
K, N = 4096, 512
packed, scale = rand_weight(fmt, K, N, seed=4)
packed[:, :64] = np.uint32(0x77777777)
x = (80.0 + np.random.default_rng(1).standard_normal(K)).astype(np.float16)
out = torch.zeros(N, dtype=torch.float16, device="cuda")
ops.GemvProgram([dict(weights=[(dev(packed), dev(scale))], x=dev(x), outs=[out])], "q4f16_0").run()
y = host(out)
y64, ab = co.gemv_f64(packed, scale, x, fmt)
assert (np.abs(y.astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all()
assert (y[:64] == 0).all()
print("EXACT-OK")
""" % (ROOT, ROOT)
    env = dict(os.environ, B200Q_PROG_TC="0")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and "EXACT-OK" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_integer_formulation_documented_env():
    """B200Q_PROG_I8=1 selects the integer (tcgen05.mma.kind::i8) formulation of the q4f16_0 program: exact s32 group
    sums over three signed base-256 digits of the activations.  Bound: tolerance_group_scaled (its activation
    quantization term); an all-q=7 column is EXACTLY zero whatever the activations (no T - C cancellation), and a
    non-finite activation poisons the outputs that depend on it."""
    code = r"""
import sys, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r + "/tests")
from mlc_llm_b200 import ops
from oracle import c_oracle as co, quant_oracle as qo
from util import dev, host, rand_weight
fmt = qo.FORMATS["q4f16_0"]
rng = np.random.default_rng(1)
for K, N, kind in ((4096, 512, "mean80"), (11008, 256, "normal"), (4096, 4096, "outlier"), (224, 136, "ragged")):
    packed, scale = rand_weight(fmt, K, N, seed=K + N)
    packed[:, :64] = np.uint32(0x77777777)
    x = rng.standard_normal(K)
    if kind == "mean80": x = 80.0 + x
    if kind == "outlier": x[::97] *= 300.0
    x = x.astype(np.float16)
    out = torch.zeros(N, dtype=torch.float16, device="cuda")
    ops.GemvProgram([dict(weights=[(dev(packed), dev(scale))], x=dev(x), outs=[out])], "q4f16_0").run()
    y = host(out)
    y64, ab, cc = co.gemv_f64_unrounded(packed, scale, x, fmt, K)
    err = np.abs(y.astype(np.float64) - y64)
    assert (err <= qo.tolerance_group_scaled(y64, ab, cc, False)).all(), (kind, float((err / qo.tolerance_group_scaled(y64, ab, cc, False)).max()))
    assert (y[:64] == 0).all(), kind
# a chain of two steps (LL-linked) and a non-finite activation
K, N = 512, 512
p1, s1 = rand_weight(fmt, K, N, seed=1); p2, s2 = rand_weight(fmt, N, 256, seed=2)
x = rng.standard_normal(K).astype(np.float16)
mid = torch.zeros(N, dtype=torch.float16, device="cuda"); out = torch.zeros(256, dtype=torch.float16, device="cuda")
prog = ops.GemvProgram([dict(weights=[(dev(p1), dev(s1))], x=dev(x), outs=[mid]), dict(weights=[(dev(p2), dev(s2))], x=mid, outs=[out])], "q4f16_0")
prog.run(); torch.cuda.synchronize()
m = host(mid)
y64, ab, cc = co.gemv_f64_unrounded(p2, s2, m, fmt, N)
assert (np.abs(host(out).astype(np.float64) - y64) <= qo.tolerance_group_scaled(y64, ab, cc, False)).all()
xn = x.copy(); xn[5] = np.inf
out2 = torch.zeros(N, dtype=torch.float16, device="cuda")
ops.GemvProgram([dict(weights=[(dev(p1), dev(s1))], x=dev(xn), outs=[out2])], "q4f16_0").run()
assert not np.isfinite(host(out2)).any()
print("I8-OK")
""" % (ROOT, ROOT)
    env = dict(os.environ, B200Q_PROG_I8="1")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and "I8-OK" in r.stdout, (r.stdout + r.stderr)[-2000:]


@pytest.mark.parametrize("use_program", [True, False])
@pytest.mark.parametrize("name", ["q4f16_0", "q8f16_0", "q3f16_0"])
def test_stack_operator_by_operator(name, use_program):
    """The whole launch sequence of a 2-layer stack (program or PDL chain): EVERY fused dequant-GEMV it ran is
    re-evaluated by the oracle from the bits it actually read and checked under the per-operator bound."""
    fmt = qo.FORMATS[name]
    shape = ModelShape("tiny-llama", "llama", 640, 1760, 2, 1000)
    st = DecodeStack(shape, name, layers=2, use_program=use_program)
    x0 = (np.random.default_rng(0).standard_normal(shape.hidden) * 0.5).astype(np.float16)
    st.x_in.copy_(torch.from_numpy(x0))
    st.capture()
    st.decode_step()
    snap = st.snapshot()
    assert len(snap) == 2 * 4 + 1

    def weights(li, nm):
        w = st.lm_head if nm == "lm_head" else st.layers[li][nm]
        return host(w.packed), host(w.scale)

    rec = stack_check.check_launches(snap, weights, fmt, group_scaled=(use_program and ops.program_uses_tensor_cores(name)))
    assert rec["ops_checked"] == 2 * 7 + 1
    assert rec["ok"], rec


def test_llama7b_two_layers_program_operator_by_operator():
    """Same, at the real Llama-7B shapes (the bench's own parity record)."""
    st = DecodeStack("llama-7b", "q4f16_0", layers=2, use_program=True)
    st.x_in.copy_((torch.randn(4096, device="cuda", generator=torch.Generator(device="cuda").manual_seed(3)) * 0.05).half())
    st.capture()
    st.decode_step()

    def weights(li, nm):
        w = st.lm_head if nm == "lm_head" else st.layers[li][nm]
        return host(w.packed), host(w.scale)

    rec = stack_check.check_launches(st.snapshot(), weights, qo.FORMATS["q4f16_0"], group_scaled=True)
    assert rec["ok"] and rec["ops_checked"] == 15, rec


def test_watchdog_turns_a_dead_peer_into_an_error():
    """A tensor-parallel program whose peer never launches: the bounded spin writes the status word and traps
    (CUDA error) instead of hanging the GPU.  Runs in a subprocess because the trap kills the context."""
    code = r"""
import sys, ctypes, numpy as np, torch
sys.path.insert(0, %r); sys.path.insert(0, %r + "/tests")
from mlc_llm_b200 import _lib, ops
from oracle import quant_oracle as qo
from util import dev, rand_weight
fmt = qo.FORMATS["q4f16_0"]
K, N = 512, 256
packed, scale = rand_weight(fmt, K, N, seed=4)
x = dev(np.ones(K, np.float16))
out = torch.zeros(N, dtype=torch.float16, device="cuda")
status = torch.zeros(1, dtype=torch.int64).pin_memory()
ops.program_watchdog(status, timeout_ms=300)
L = _lib.lib()
arr = (_lib.GemvStep * 1)()
arr[0].n_ops = 1; arr[0].x = x.data_ptr(); arr[0].x_mul = None; arr[0].K = K
pk, sc = dev(packed), dev(scale)
arr[0].ops[0] = _lib.GemvOp(pk.data_ptr(), sc.data_ptr(), None, out.data_ptr(), N, 0)
n = int(L.b200q_gemv_program_bytes(1))
blob = np.zeros(n // 8 + 1, np.uint64)
mask = np.array([1], np.uint8)
_lib.check(L.b200q_gemv_program_build_tp(arr, 1, 0, mask.ctypes.data, 0, 2, blob.ctypes.data, n))
dblob = torch.from_numpy(blob.view(np.uint8)[:n].copy()).cuda()
wsb = int(L.b200q_gemv_program_workspace_bytes(blob.ctypes.data))
ws0 = torch.zeros(wsb, dtype=torch.uint8, device="cuda"); ws1 = torch.zeros(wsb, dtype=torch.uint8, device="cuda")
peers = (ctypes.c_void_p * 2)(ws0.data_ptr(), ws1.data_ptr())   # rank 1 never runs
_lib.check(L.b200q_gemv_program_run_tp(blob.ctypes.data, dblob.data_ptr(), ws0.data_ptr(), wsb, peers, torch.cuda.current_stream().cuda_stream))
try:
    torch.cuda.synchronize()
    print("NO-ERROR")
except Exception as e:
    print("CUDA-ERROR", type(e).__name__)
v = int(status[0].item())
print("STATUS", hex(v & 0xFFFF), (v >> 16) & 0xFFFF, v >> 32)
""" % (ROOT, ROOT)
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    out = r.stdout + r.stderr
    assert "CUDA-ERROR" in out and "STATUS 0xdead 4 0" in out, out[-3000:]
