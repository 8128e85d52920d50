"""GPU parity tests of the prefill path: the tcgen05 dequant-GEMM (b200q_gemm) and the dense NT GEMM
(b200q_nt_matmul) against the oracle (checker only).

Bar: the dequantized weights fed to the tensor cores are the bit-exact reference
weights (checked here through X = identity: Y == decode(W) bit for bit, the f16 x f16 products and the
f32 sums being exact), and Y is within  2^-10 |y64| + 2^-14 sum|x w| + 2^-24  of the f64 evaluation
on the same dequantized weights.  The reference's own f16-accumulating NT_matmul is compared too.
"""
import numpy as np
import pytest
import torch

from mlc_llm_b200 import dispatch, ops
from oracle import c_oracle as co
from oracle import quant_oracle as qo
from util import beq, dev, host, rand_weight

pytestmark = pytest.mark.gpu

F16 = ["q4f16_0", "q3f16_0", "q8f16_0"]


def _check(Y, packed, scale, X, fmt, what, out_f32=False):
    y64, ab = qo.gemm_f64(packed, scale, X, fmt)
    tol = qo.tolerance(y64, ab, out_f32)
    err = np.abs(Y.astype(np.float64) - y64)
    worst = float((err / tol).max())
    assert worst <= 1.0, f"{what}: error/tolerance = {worst:.3f} (max abs err {err.max():.3e})"


@pytest.mark.parametrize("name", F16)
def test_gemm_identity_reproduces_decode_bit_exact(name):
    """X = I (K x K): Y[k, n] must be the dequantized weight itself - every k row, every code."""
    fmt = qo.FORMATS[name]
    K, N = 320, 512  # 320 = lcm-friendly for all three: 10 groups of 32, 8 groups of 40
    packed, scale = rand_weight(fmt, K, N, seed=77)
    X = np.eye(K, dtype=np.float16)
    Y = host(ops.gemm(dev(packed), dev(scale), dev(X), name))
    assert beq(Y, co.decode(packed, scale, K, fmt))


@pytest.mark.parametrize("name", F16)
@pytest.mark.parametrize("M,K,N", [(128, 256, 256), (256, 2560, 512), (2048, 2560, 2560), (300, 640, 1000),
                                   (1, 4096, 4096), (77, 1280, 264), (512, 10240, 256)])
def test_gemm_tolerance_shapes(name, M, K, N):
    fmt = qo.FORMATS[name]
    packed, scale = rand_weight(fmt, K, N, seed=M + K + N)
    X = qo.synth_act((M, K), fmt, seed=M)
    Y = host(ops.gemm(dev(packed), dev(scale), dev(X), name))
    assert Y.shape == (M, N) and Y.dtype == np.float16
    if M * K * N > 2**31:  # keep the f64 check in seconds: sample rows
        rows = np.random.default_rng(0).choice(M, 64, replace=False)
        _check(Y[rows], packed, scale, X[rows], fmt, f"{name} {M}x{K}x{N}")
    else:
        _check(Y, packed, scale, X, fmt, f"{name} {M}x{K}x{N}")


@pytest.mark.parametrize("name", F16)
def test_gemm_encoded_weights_ragged_k(name):
    """Encoder-produced weights with K not a multiple of the group / of the stage depth."""
    fmt = qo.FORMATS[name]
    for K in (200, 1000, 4096):
        N, M = 264, 130
        W = qo.synth_weight(N, K, fmt, seed=K)
        packed, scale = qo.encode_sym(W, fmt)
        X = qo.synth_act((M, K), fmt, seed=3)
        Y = host(ops.gemm(dev(packed), dev(scale), dev(X), name))
        y64 = X.astype(np.float64) @ qo.decode_sym(packed, scale, K, fmt).astype(np.float64)
        ab = np.abs(X.astype(np.float64)) @ np.abs(qo.decode_sym(packed, scale, K, fmt).astype(np.float64))
        err = np.abs(Y.astype(np.float64) - y64)
        assert (err <= qo.tolerance(y64, ab, False)).all(), (name, K, float((err / qo.tolerance(y64, ab, False)).max()))


def test_gemm_row_equals_gemv():
    """Row m of the prefill GEMM and the decode GEMV on the same activation are two f32 sums of the
    same exact products in different orders: both inside the tolerance, hence within 2x of each other."""
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    K, N, M = 4096, 1024, 64
    packed, scale = rand_weight(fmt, K, N, seed=5)
    X = qo.synth_act((M, K), fmt, seed=6)
    p, s = dev(packed), dev(scale)
    Y = host(ops.gemm(p, s, dev(X), name))
    y64, ab = qo.gemm_f64(packed, scale, X, fmt)
    tol = qo.tolerance(y64, ab, False)
    for m in (0, 17, 63):
        y = host(ops.gemv(p, s, dev(X[m]), name)).reshape(-1)
        assert (np.abs(Y[m].astype(np.float64) - y.astype(np.float64)) <= 2 * tol[m]).all()
        assert (np.abs(Y[m].view(np.int16).astype(np.int32) - y.view(np.int16).astype(np.int32)) <= 1).mean() > 0.99


@pytest.mark.parametrize("name", F16)
def test_gemm_epilogues_op_for_op(name):
    fmt = qo.FORMATS[name]
    M, K, N = 200, 640, 520
    packed, scale = rand_weight(fmt, K, N, seed=31)
    X = qo.synth_act((M, K), fmt, seed=32)
    bias = qo.synth_act(N, fmt, seed=33)
    r1 = qo.synth_act((M, N), fmt, seed=34)
    r2 = qo.synth_act((M, N), fmt, seed=35)
    p, s, xd = dev(packed), dev(scale), dev(X)
    y0 = host(ops.gemm(p, s, xd, name))
    for ep, e0 in (("add", r1), ("multiply", r1), ("bias", bias)):
        y = host(ops.gemm(p, s, xd, name, ep, dev(e0)))
        assert beq(y, qo.apply_epilogue(y0, qo.EPILOGUES[ep], e0)), ep
    y = host(ops.gemm(p, s, xd, name, "bias_add2", dev(bias), dev(r1), dev(r2)))
    assert beq(y, qo.apply_epilogue(y0, qo.EP_BIAS_ADD2, bias, r1, r2))
    for ep in ("silu", "bias_gelu"):
        y = host(ops.gemm(p, s, xd, name, ep, dev(bias) if ep == "bias_gelu" else None))
        e = qo.apply_epilogue(y0, qo.EPILOGUES[ep], bias)
        ulp = np.abs(y.view(np.int16).astype(np.int32) - e.view(np.int16).astype(np.int32))
        assert ulp.max() <= 1, (ep, ulp.max())
    y32 = host(ops.gemm(p, s, xd, name, "cast_f32"))
    assert y32.dtype == np.float32 and beq(y32.astype(np.float16), y0)
    _check(y32, packed, scale, X, fmt, "cast_f32", out_f32=True)


@pytest.mark.parametrize("M,K,N", [(128, 128, 256), (200, 2560, 520), (1024, 1024, 1024), (5, 64, 8)])
def test_nt_matmul_dense(M, K, N):
    """NT_matmul on a dense f16 weight T[N, K]: within tolerance of f64, and at least as close to f64
    as the reference's f16-accumulating loop."""
    rng = np.random.default_rng(M + N)
    X = rng.standard_normal((M, K)).astype(np.float16)
    T = (rng.standard_normal((N, K)) * 0.02).astype(np.float16)
    Y = host(ops.nt_matmul(dev(X), dev(T)))
    y64, ab = co.nt_matmul_f64_f16(X, T)
    err = np.abs(Y.astype(np.float64) - y64)
    assert (err <= qo.tolerance(y64, ab, False)).all()
    if M * K * N <= 2**27:
        yref = co.nt_matmul_ref_f16(X, T)
        assert err.max() <= np.abs(yref.astype(np.float64) - y64).max() + 1e-12


def test_two_kernel_prefill_equals_fused():
    """The reference-shaped route (decode to dense [N, K], then NT_matmul) and the fused dequant-GEMM
    see the same weights and the same k order inside a 16-k MMA step: results agree to 1 f16 ulp."""
    name = "q4f16_0"
    fmt = qo.FORMATS[name]
    M, K, N = 256, 1024, 512
    packed, scale = rand_weight(fmt, K, N, seed=8)
    X = qo.synth_act((M, K), fmt, seed=9)
    p, s, xd = dev(packed), dev(scale), dev(X)
    T = ops.dequantize(p, s, K, name, transpose_output=True)
    y2 = host(ops.nt_matmul(xd, T))
    y1 = host(ops.gemm(p, s, xd, name, streamk=False))
    assert beq(y1, y2)  # same weights, same tile shape, same k order (whole-tile schedule): identical f32 sums
    y3 = host(ops.gemm(p, s, xd, name, streamk="force"))  # stream-K cuts the k range of a tile: f32 rounding of the partial sums only
    d = np.abs(y3.astype(np.float64) - y2.astype(np.float64))
    assert (d <= 2.0 ** -9 * np.abs(y2.astype(np.float64)) + 2.0 ** -14).all()


def test_gemm_dispatch_lookup_and_errors():
    fmt = qo.FORMATS["q3f16_0"]
    M, K, N = 64, 2560, 256
    packed, scale = rand_weight(fmt, K, N, seed=1)
    X = qo.synth_act((1, M, K), fmt, seed=2)
    f = dispatch.lookup("q3f16_0", K, N, "none", m_is_one=False)
    assert f is not None and "NT_matmul" in f.__name__
    out = torch.empty((1, M, N), dtype=torch.float16, device="cuda")
    f(dev(packed), dev(scale), dev(X), out)
    _check(host(out)[0], packed, scale, X[0], fmt, "dispatch q3")
    assert dispatch.lookup("q4f32_0", K, N, "none", m_is_one=False) is None  # f32 prefill: a miss, like the reference
    with pytest.raises(Exception):
        ops.gemm(dev(packed), dev(scale), dev(X[0, :, :-4].copy()), "q3f16_0")  # K mismatch


@pytest.mark.parametrize("tile", ["single", "pair128", "pair256"])
@pytest.mark.parametrize("name", F16)
def test_gemm_every_tile_kernel(name, tile):
    """The single-CTA kernel and both CTA-pair (cta_group::2) tile shapes on the same inputs: X = I reproduces the
    decode bit for bit through each, ragged M / N / K tails included, and all three agree with the oracle."""
    fmt = qo.FORMATS[name]
    K, N = 320, 520
    packed, scale = rand_weight(fmt, K, N, seed=71)
    X = np.eye(K, dtype=np.float16)
    Y = host(ops.gemm(dev(packed), dev(scale), dev(X), name, tile=tile))
    assert beq(Y, co.decode(packed, scale, K, fmt)), tile
    for M, K, N in ((300, 640, 1000), (2048, 2560, 2560), (130, 1000, 264), (512, 10240, 256), (257, 320, 136)):
        if fmt.nbit == 3 and K % 40:
            K = K // 40 * 40
        packed, scale = rand_weight(fmt, K, N, seed=M + N)
        X = qo.synth_act((M, K), fmt, seed=K)
        bias = qo.synth_act(N, fmt, seed=2)
        Y = host(ops.gemm(dev(packed), dev(scale), dev(X), name, "bias", dev(bias), tile=tile))
        Y0 = host(ops.gemm(dev(packed), dev(scale), dev(X), name, tile=tile))
        assert beq(Y, qo.apply_epilogue(Y0, qo.EP_BIAS, np.broadcast_to(bias, (M, N)))), (tile, M, K, N)
        rows = np.random.default_rng(0).choice(M, min(M, 48), replace=False)
        _check(Y0[rows], packed, scale, X[rows], fmt, f"{name} {tile} {M}x{K}x{N}")


@pytest.mark.parametrize("name", F16 + ["q4f16_asym"])
@pytest.mark.parametrize("M,K,N", [(2048, 2560, 2560), (2048, 10240, 2560), (300, 640, 1000), (512, 1280, 256), (256, 320, 264),
                                   (1030, 2560, 520)])
def test_gemm_streamk_schedule(name, M, K, N):
    """b200q_gemm_ws: the stream-K schedule of the CTA-pair kernel (a tile's k range cut over several pairs, f32
    partial tiles parked in the workspace, the finisher adds them in pair order) against the oracle, against the
    whole-tile schedule, and launch after launch on the same workspace (flags lowered by the finisher)."""
    from mlc_llm_b200.quant_spec import SPECS
    spec = SPECS[name]
    if spec.nbit == 3 and K % 40:
        K = K // 40 * 40
    rng = np.random.default_rng(M + K + N)
    if spec.sym:
        fmt = qo.FORMATS[name]
        packed, scale = rand_weight(fmt, K, N, seed=M + N + 1)
        zmin = None
    else:
        packed = rng.integers(0, 2**32, size=(spec.n_words(K), N), dtype=np.uint64).astype(np.uint32)
        scale = (rng.random((spec.n_groups(K), N), dtype=np.float32) * 0.01).astype(np.float16)
        zmin = (-(rng.random((spec.n_groups(K), N), dtype=np.float32)) * 0.07).astype(np.float16)
    X = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    p, s, x = dev(packed), dev(scale), dev(X)
    z = dev(zmin) if zmin is not None else None
    ws = torch.zeros(int(ops._lib.lib().b200q_gemm_workspace_bytes()), dtype=torch.uint8, device="cuda")
    whole = host(ops.gemm(p, s, x, name, zmin=z, ws=ws, streamk=False, tile="pair256"))
    assert beq(whole, host(ops.gemm(p, s, x, name, zmin=z, ws=False, tile="pair256")))  # NO_STREAMK == no workspace
    outs = [host(ops.gemm(p, s, x, name, zmin=z, ws=ws, tile="pair256", streamk="force")) for _ in range(3)]
    torch.cuda.synchronize()
    assert int(ws[:16384].max()) == 0, "the finisher lowers every flag it consumed"
    assert beq(outs[0], outs[1]) and beq(outs[0], outs[2]), "deterministic, launch after launch"
    # both schedules within the oracle's tolerance of the f64 result; they differ from one another by f32 rounding only
    rows = rng.choice(M, min(M, 40), replace=False)
    if spec.sym:
        _check(outs[0][rows], packed, scale, X[rows], qo.FORMATS[name], f"stream-K {name} {M}x{K}x{N}")
    else:
        w = qo.decode_asym(packed, scale, zmin, K, qo.ASYM_FORMATS[name]).astype(np.float64)
        y64 = X[rows].astype(np.float64) @ w
        ab = np.abs(X[rows].astype(np.float64)) @ np.abs(w)
        assert (np.abs(outs[0][rows].astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all()
    if spec.sym:
        y64, ab = qo.gemm_f64(packed, scale, X[rows], qo.FORMATS[name])
    d = np.abs(outs[0][rows].astype(np.float64) - whole[rows].astype(np.float64))
    assert (d <= 2 * qo.tolerance(y64, ab, False)).all(), float(d.max())
    # epilogue through the finisher
    bias = (rng.standard_normal(N) * 0.1).astype(np.float16)
    yb = host(ops.gemm(p, s, x, name, "bias", dev(bias), zmin=z, ws=ws, tile="pair256", streamk="force"))
    assert beq(yb, qo.apply_epilogue(outs[0], qo.EP_BIAS, np.broadcast_to(bias, (M, N))))


@pytest.mark.parametrize("name", F16)
@pytest.mark.parametrize("M,K,Ns", [(2048, 2560, (2560, 2560, 2560)), (300, 640, (264, 520, 136)), (512, 1280, (1280, 256)),
                                    (100, 320, (264, 264, 264))])
def test_gemm_multi_equals_single_launches(name, M, K, Ns):
    """b200q_gemm_multi: q | k | v in one launch of the pair kernel (one tile space over the three weights) - the
    whole-tile schedule reproduces the three single launches bit for bit, the stream-K schedule stays inside the
    oracle's bound; per-weight bias epilogues; M <= 128 falls back to one launch per weight."""
    fmt = qo.FORMATS[name]
    if fmt.nbit == 3 and K % 40:
        K = K // 40 * 40
    rng = np.random.default_rng(M + K)
    X = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    xd = dev(X)
    ws, host_w, biases = [], [], []
    for i, N in enumerate(Ns):
        packed, scale = rand_weight(fmt, K, N, seed=10 * i + N)
        host_w.append((packed, scale))
        ws.append((dev(packed), dev(scale)))
        biases.append(dev((rng.standard_normal(N) * 0.1).astype(np.float16)))
    eps = ["bias"] * len(Ns)
    single = [host(ops.gemm(w[0], w[1], xd, name, "bias", b, ws=False)) for w, b in zip(ws, biases)]
    multi = [host(y) for y in ops.gemm_multi(ws, xd, name, eps, biases, streamk=False)]
    for a, b in zip(single, multi):
        assert beq(a, b)
    sk = [host(y) for y in ops.gemm_multi(ws, xd, name, eps, biases, streamk="force")]
    sk2 = [host(y) for y in ops.gemm_multi(ws, xd, name, eps, biases, streamk="force")]
    rows = rng.choice(M, min(M, 24), replace=False)
    for (packed, scale), y, y2, bias, ref in zip(host_w, sk, sk2, biases, single):
        assert beq(y, y2)
        y0 = host(ops.gemm(dev(packed), dev(scale), xd, name, streamk="force")) if M > 128 else None
        y64, ab = qo.gemm_f64(packed, scale, X[rows], fmt)
        d = np.abs(y[rows].astype(np.float64) - ref[rows].astype(np.float64))
        assert (d <= 2 * qo.tolerance(y64, ab, False) + 2.0 ** -10 * np.abs(ref[rows].astype(np.float64))).all()


@pytest.mark.parametrize("name", F16)
def test_gemm_gelu_table_equals_erff(name):
    """bias + GELU on 256-wide pair tiles reads erf from a shared-memory table of f16(erff(t)) for every f16 t in [0, 4)
    (csrc/gemm.cu, kGelu instantiation); the other tile kernels call erff(): identical bits, also for |t| >= 4, tiny t
    and both signs."""
    fmt = qo.FORMATS[name]
    M, K, N = 300, 640, 520
    if fmt.nbit == 3:
        K = 640 // 40 * 40
    packed, scale = rand_weight(fmt, K, N, seed=5)
    scale = (scale.astype(np.float32) * 40).astype(np.float16)  # |t| up to ~10
    rng = np.random.default_rng(6)
    X = rng.standard_normal((M, K)).astype(np.float16)
    X[:40] *= np.float16(1e-3)  # rows with tiny pre-activations
    bias = (rng.standard_normal(N) * 0.5).astype(np.float16)
    p, s, xd, b = dev(packed), dev(scale), dev(X), dev(bias)
    lut = host(ops.gemm(p, s, xd, name, "bias_gelu", b, tile="pair256", ws=False))
    for other in ("pair128", "single"):
        assert beq(lut, host(ops.gemm(p, s, xd, name, "bias_gelu", b, tile=other, ws=False))), other
    y0 = host(ops.gemm(p, s, xd, name, ws=False))
    e = qo.apply_epilogue(y0, qo.EP_BIAS_GELU, np.broadcast_to(bias, (M, N)))
    ulp = np.abs(lut.view(np.int16).astype(np.int32) - e.view(np.int16).astype(np.int32))
    assert ulp.max() <= 1
    t = (y0.astype(np.float32) + bias) * 0.7071
    assert (np.abs(t) >= 4).any() and (np.abs(t) < 1e-2).any()


@pytest.mark.parametrize("M,K,N", [(2048, 2560, 2560), (300, 640, 520)])
def test_nt_matmul_streamk(M, K, N):
    """dense NT_matmul through the same pair kernel with the stream-K workspace: deterministic, inside the bound"""
    rng = np.random.default_rng(M + N)
    X = rng.standard_normal((M, K)).astype(np.float16)
    T = (rng.standard_normal((N, K)) * 0.02).astype(np.float16)
    xd, td = dev(X), dev(T)
    whole = host(ops.nt_matmul(xd, td, streamk=False))
    assert beq(whole, host(ops.nt_matmul(xd, td, ws=False)))
    a = host(ops.nt_matmul(xd, td, streamk="force"))
    assert beq(a, host(ops.nt_matmul(xd, td, streamk="force")))
    rows = rng.choice(M, 32, replace=False)
    y64, ab = co.nt_matmul_f64_f16(X[rows], T)
    assert (np.abs(a[rows].astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all()
    assert (np.abs(whole[rows].astype(np.float64) - y64) <= qo.tolerance(y64, ab, False)).all()


def test_gemm_multi_four_weights_with_gelu():
    """q | k | v | h_to_4h of a GPT-NeoX layer in one launch (the GELU-table instantiation with four weights): each
    output equals the single launch of its weight under the whole-tile schedule; stream-K stays inside the bound."""
    name = "q3f16_0"
    fmt = qo.FORMATS[name]
    M, K, Ns = 520, 640, (264, 264, 264, 1032)
    rng = np.random.default_rng(3)
    X = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    xd = dev(X)
    ws, hw, bs = [], [], []
    for i, N in enumerate(Ns):
        packed, scale = rand_weight(fmt, K, N, seed=40 + i)
        hw.append((packed, scale)); ws.append((dev(packed), dev(scale)))
        bs.append(dev((rng.standard_normal(N) * 0.3).astype(np.float16)))
    eps = ["bias", "bias", "bias", "bias_gelu"]
    single = [host(ops.gemm(w[0], w[1], xd, name, e, b, ws=False)) for w, e, b in zip(ws, eps, bs)]
    multi = [host(y) for y in ops.gemm_multi(ws, xd, name, eps, bs, streamk=False)]
    for a, b in zip(single, multi):
        assert beq(a, b)
    sk = [host(y) for y in ops.gemm_multi(ws, xd, name, eps, bs, streamk="force")]
    for a, b in zip(single, sk):
        d = np.abs(a.astype(np.float64) - b.astype(np.float64))
        assert (d <= 2.0 ** -8 * np.abs(a.astype(np.float64)) + 2.0 ** -8).all()
    # four weights without a GELU tail: two launches behind the same call, same results
    multi2 = [host(y) for y in ops.gemm_multi(ws, xd, name, ["bias"] * 4, bs, streamk=False)]
    for (packed, scale), y, b in zip(hw, multi2, bs):
        assert beq(y, host(ops.gemm(dev(packed), dev(scale), xd, name, "bias", b, ws=False)))
