"""The reference's OWN hand-scheduled GPU kernels (TVM-generated CUDA of mlc_llm/dispatch PrimFuncs, compiled for
sm_100a by oracle/build_ref_cuda.py) run on the B200: their outputs pin the oracle once more - the fused GEMVs are
the reference's dtype-accumulating chains (compared with the oracle's reference-semantics GEMV inside the
reference's own acceptance band, tests/debug/compare_lib.py:26-27), the decode kernel is bit-exact - and the same
inputs through this library's kernels land inside the f64 tolerance.  Skipped when oracle/_ref/cuda is absent."""
import numpy as np
import pytest
import torch

from mlc_llm_b200 import ops
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from oracle import ref_cuda
from util import beq, dev, host, rand_weight

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref_cuda.available(), reason="oracle/_ref/cuda not built")]


def test_reference_gemv_q4f16_0_kernel_matches_oracle_and_ours():
    fmt = qo.FORMATS["q4f16_0"]
    K = N = 4096
    packed, scale = rand_weight(fmt, K, N, seed=11)
    x = qo.synth_act(K, fmt, seed=12)
    out = torch.zeros((1, 1, N), dtype=torch.float16, device="cuda")
    k = ref_cuda.RefKernel("gemv_q4f16_0_4096x4096")
    k(dev(packed), dev(scale), dev(x.reshape(1, 1, K)), out)  # fused_decode4_matmul3(lv1587, lv1588, lv1583, out)
    torch.cuda.synchronize()
    y_ref_gpu = host(out).reshape(-1)
    y_orc = co.gemv_ref(packed, scale, x, fmt)  # the reference's f16 chain, k ascending
    y64, ab = co.gemv_f64(packed, scale, x, fmt)
    # the scheduled kernel splits K (different f16 summation order than the unscheduled chain): same band
    assert np.abs(y_ref_gpu.astype(np.float64) - y_orc.astype(np.float64)).max() <= 1e-2 + 1e-3 * np.abs(y_orc).max() + 2 * np.abs(y_orc.astype(np.float64) - y64).max()
    ours = host(ops.gemv(dev(packed), dev(scale), dev(x), "q4f16_0")).reshape(-1)
    err = np.abs(ours.astype(np.float64) - y64)
    assert (err <= qo.tolerance(y64, ab, False)).all()
    assert err.max() <= np.abs(y_ref_gpu.astype(np.float64) - y64).max()  # closer to the truth than the reference kernel


def test_reference_decode_kernel_bit_exact_with_oracle_and_ours():
    """decode4_fp16 (asymmetric, transposed output) scheduled by decode_sch_func."""
    fmt = qo.ASYM_FORMATS["q4f16_asym"]
    K = N = 4096
    rng = np.random.default_rng(3)
    packed = rng.integers(0, 2**32, size=(K // 8, N), dtype=np.uint64).astype(np.uint32)
    sc = (rng.random((K // 32, N), dtype=np.float32) * 0.01).astype(np.float16)
    mn = (-rng.random((K // 32, N), dtype=np.float32) * 0.08).astype(np.float16)
    out = torch.zeros((N, K), dtype=torch.float16, device="cuda")
    k = ref_cuda.RefKernel("decode_q4f16_asym_4096x4096")
    k(dev(packed), dev(sc), dev(mn), out)
    torch.cuda.synchronize()
    w_ref = host(out)
    w_orc = co.decode_asym(packed, sc, mn, K, fmt, transpose_output=True)
    w_ours = host(ops.dequantize(dev(packed), dev(sc), K, "q4f16_asym", True, zmin=dev(mn)))
    assert beq(w_ours, w_orc)
    # nvcc contracts  q * s + m  into one fma in the reference's generated CUDA (the codegen ambiguity SURVEY.md 8c
    # names; the reference's CPU builds - and therefore the oracle and this library - round the product first):
    # the two differ by at most the rounding of the product (2^-11 * |q s| <= 2^-11 * 15 * 0.01) plus one ulp of the
    # result that this can flip, and mostly agree
    d = np.abs(w_ref.astype(np.float64) - w_orc.astype(np.float64))
    assert (d <= 2.0**-11 * 15 * 0.01 + 2.0**-10 * np.abs(w_orc.astype(np.float64))).all(), d.max()
    assert (w_ref == w_orc).mean() > 0.5


def test_reference_nt_matmul_kernel_vs_ours():
    M, K, N = 2048, 4096, 4096
    rng = np.random.default_rng(5)
    X = (rng.standard_normal((1, M, K)) * 0.5).astype(np.float16)
    T = (rng.standard_normal((N, K)) * 0.02).astype(np.float16)
    out = torch.zeros((1, M, N), dtype=torch.float16, device="cuda")
    k = ref_cuda.RefKernel("nt_matmul_f16_4096x4096")
    k(dev(X), dev(T), out)
    torch.cuda.synchronize()
    y_ref = host(out)[0]
    ours = host(ops.nt_matmul(dev(X[0]), dev(T)))
    rows = np.random.default_rng(0).choice(M, 32, replace=False)
    y64 = X[0][rows].astype(np.float64) @ T.astype(np.float64).T
    e_ours, e_ref = np.abs(ours[rows] - y64).max(), np.abs(y_ref[rows] - y64).max()
    assert e_ours <= e_ref + 1e-12 and e_ref <= 0.1 * max(1.0, np.abs(y64).max())
