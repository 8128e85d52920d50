"""Generate tests/golden/*.npz by running the REFERENCE's own code in this container.

TEST INFRASTRUCTURE ONLY.  Needs /root/reference (read-only) and the TVM 0.23.dev0 that the
`tilelang` wheel bundles (there is no stand-alone tvm here).  Neither exists on
the GPU box, so the vectors this script writes are committed and the tests read only those.

What is executed (unmodified reference code, loaded with importlib):
  * mlc_llm/transform/quantization.py  encoding_func / decoding_func / decoding_after_taking_func
    -> te.create_prim_func -> tvm.compile(target="c") -> g++ (-Dhalf=_Float16 -fexcess-precision=16)
  * mlc_llm/dispatch/dispatch_tir_operator_adreno.py   the unscheduled q4f16_0 fused GEMV PrimFuncs
    (fused_decode4_matmul3, ..._add1, ..._multiply1, ..._silu1, ..._cast2) at their Llama-7B shapes
  * mlc_llm/dispatch/llama/main.py   the unscheduled int3 "before" fused GEMVs

Harness-side shims (the reference is untouched):  tvm.nd alias (API removed in 0.23) and C lowerings
for tir.fabs / tir.round (-> fabsf / roundf), which the bundled TVM's C host lacks.

Usage:  PYTHONDONTWRITEBYTECODE=1 python oracle/gen_golden.py [--skip-fused]
"""
from __future__ import annotations

import argparse
import hashlib
import importlib.util
import os
import sys
import tempfile
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF = "/root/reference"
sys.path.insert(0, ROOT)

from oracle import quant_oracle as qo  # noqa: E402

CC_OPTS = ["-Dhalf=_Float16", "-fexcess-precision=16", "-O2", "-include", "math.h"]


def _load_tvm():
    import tilelang  # noqa: F401  (puts its bundled tvm on sys.path)
    import tvm
    from tvm import tir

    tvm.nd = types.SimpleNamespace(NDArray=tvm.runtime.Tensor, array=tvm.runtime.tensor)

    def _ext(name):
        def f(op):
            x = op.args[0]
            if x.dtype == "float32":
                return tir.call_pure_extern("float32", name, x)
            return tir.Cast(x.dtype, tir.call_pure_extern("float32", name, tir.Cast("float32", x)))

        return f

    for op, c in [("tir.fabs", "fabsf"), ("tir.round", "roundf"), ("tir.erf", "erff"), ("tir.exp", "expf")]:
        try:
            tvm.ir.register_intrin_lowering(op, target="c", f=_ext(c), level=99)
        except Exception as e:  # already registered
            print("lowering", op, e)
    return tvm


def _load_ref_quant():
    spec = importlib.util.spec_from_file_location("refq", f"{REF}/mlc_llm/transform/quantization.py")
    refq = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(refq)
    return refq


_tmp = tempfile.mkdtemp(prefix="b200q_golden_")
_cnt = [0]


def build_run(tvm, prim_func, args_np):
    """Compile a PrimFunc with TVM's C backend and run it on numpy arrays (outputs in place)."""
    _cnt[0] += 1
    path = os.path.join(_tmp, f"f{_cnt[0]}.so")
    lib = tvm.compile(prim_func.with_attr("global_symbol", "main"), target="c")
    lib.export_library(path, options=CC_OPTS)
    mod = tvm.runtime.load_module(path)
    targs = [tvm.runtime.tensor(a) for a in args_np]
    t0 = time.time()
    mod["main"](*targs)
    dt = time.time() - t0
    return [t.numpy() for t in targs], dt


from oracle.golden_inputs import sha, synth_fused_inputs  # noqa: E402


def gen_encode_decode(tvm, refq):
    from tvm import te

    out = {}
    cases = [
        # fmt name, N, K
        ("q4f16_0", 24, 256),
        ("q4f16_0", 16, 4096),
        ("q4f32_0", 24, 256),
        ("q4f32_0", 8, 2560),
        ("q3f16_0", 24, 320),
        ("q3f16_0", 24, 100),  # ragged: K % 40 != 0 and K % 10 == 0
        ("q3f16_0", 8, 4096),  # the Llama case: padded to 4120 (main.py:4097  u32[412, N])
        ("q8f16_0", 24, 256),
        ("q8f16_0", 8, 5120),
    ]
    for ci, (name, N, K) in enumerate(cases):
        fmt = qo.FORMATS[name]
        mode = f"int{fmt.nbit}"
        rng = np.random.default_rng(1000 + ci)
        W = (rng.standard_normal((N, K), dtype=np.float32) * 0.02).astype(fmt.np_dtype)
        # plant edge values: an all-zero group (scale floor 1e-4), exact halves, a huge outlier
        W[0, : fmt.group] = 0
        W[1, 0] = fmt.np_dtype(0.5)
        W[2, 1] = fmt.np_dtype(-0.25)
        R, G = fmt.n_words(K), fmt.n_groups(K)

        enc = refq.encoding_func(True, fmt.group, fmt.nbit, mode, 32, transpose=True, dtype=fmt.dtype)
        Wt = te.placeholder((N, K), fmt.dtype, name="W")
        packed_t, scale_t = enc(Wt)
        f = te.create_prim_func([Wt, packed_t, scale_t])
        (_, packed, scale), _ = build_run(
            tvm, f, [W, np.zeros((R, N), np.uint32), np.zeros((G, N), fmt.np_dtype)]
        )

        res = {"W": W, "packed": packed, "scale": scale}
        for tr in (False, True):
            dec = refq.decoding_func(
                True, fmt.group, fmt.nbit, mode, 32, K, data_transposed=True, transpose_output=tr, dtype=fmt.dtype
            )
            D = te.placeholder((R, N), "uint32", name="D")
            S = te.placeholder((G, N), fmt.dtype, name="S")
            f = te.create_prim_func([D, S, dec(D, S)])
            shape = (N, K) if tr else (K, N)
            (_, _, w), _ = build_run(tvm, f, [packed, scale, np.zeros(shape, fmt.np_dtype)])
            res["w_t" if tr else "w"] = w

        # decode of arbitrary words: all codes incl. those the encoder never emits (15 / 7 / 255)
        rw = rng.integers(0, 2**32, size=(R, N), dtype=np.uint64).astype(np.uint32)
        rs = (rng.random((G, N), dtype=np.float32) * 0.01).astype(fmt.np_dtype)
        dec = refq.decoding_func(
            True, fmt.group, fmt.nbit, mode, 32, K, data_transposed=True, transpose_output=False, dtype=fmt.dtype
        )
        D = te.placeholder((R, N), "uint32", name="D")
        S = te.placeholder((G, N), fmt.dtype, name="S")
        f = te.create_prim_func([D, S, dec(D, S)])
        (_, _, w), _ = build_run(tvm, f, [rw, rs, np.zeros((K, N), fmt.np_dtype)])
        res.update(rand_packed=rw, rand_scale=rs, rand_w=w)

        # embedding path (transpose=False layout + take_decode), only where K % group == 0
        if K % fmt.group == 0 and N >= 16:
            enc = refq.encoding_func(True, fmt.group, fmt.nbit, mode, 32, transpose=False, dtype=fmt.dtype)
            packed_t, scale_t = enc(Wt)
            f = te.create_prim_func([Wt, packed_t, scale_t])
            (_, p_rm, s_rm), _ = build_run(
                tvm, f, [W, np.zeros((N, R), np.uint32), np.zeros((N, G), fmt.np_dtype)]
            )
            ids = np.array([3, 0, N - 1, 3, 7], dtype=np.int32)
            tk = refq.decoding_after_taking_func(True, fmt.group, fmt.nbit, mode, 32, K, dtype=fmt.dtype)
            D = te.placeholder((N, R), "uint32", name="D")
            S = te.placeholder((N, G), fmt.dtype, name="S")
            I = te.placeholder((len(ids),), "int32", name="I")
            f = te.create_prim_func([D, S, I, tk(D, S, I)])
            (_, _, _, emb), _ = build_run(tvm, f, [p_rm, s_rm, ids, np.zeros((len(ids), K), fmt.np_dtype)])
            res.update(packed_rm=p_rm, scale_rm=s_rm, take_ids=ids, take_out=emb)

        for k_, v_ in res.items():
            out[f"{name}_N{N}_K{K}/{k_}"] = v_
        print(f"encode/decode {name} N={N} K={K}: ok")
    np.savez_compressed(os.path.join(GOLDEN, "encode_decode.npz"), **out)


def gen_fused(tvm):
    sys.path.insert(0, REF)
    import mlc_llm  # noqa: F401
    from mlc_llm.dispatch import dispatch_tir_operator_adreno as ad

    out = {}
    fmt = qo.FORMATS["q4f16_0"]
    # (prim func, K, N, epilogue name, out dtype)
    cases = [
        ("fused_decode4_matmul3", 4096, 4096, "none", np.float16),
        ("fused_decode6_fused_matmul7_add1", 11008, 4096, "add", np.float16),
        ("fused_decode5_fused_matmul6_multiply1", 4096, 11008, "multiply", np.float16),
        ("fused_decode5_fused_matmul6_silu1", 4096, 11008, "silu", np.float16),
        ("fused_decode4_fused_matmul4_add1", 4096, 4096, "add", np.float16),
        ("fused_decode3_fused_matmul1_cast2", 4096, 32000, "cast_f32", np.float32),
    ]
    for ci, (fname, K, N, ep, odt) in enumerate(cases):
        seed = 7000 + ci
        packed, scale, x, epv = synth_fused_inputs(fmt, K, N, seed)
        args = [packed, scale, x.reshape(1, 1, K)]
        if ep in ("add", "multiply"):
            args.append(epv.reshape(1, 1, N))
        args.append(np.zeros((1, 1, N), odt))
        res, dt = build_run(tvm, getattr(ad, fname), args)
        y = res[-1].reshape(N)
        key = f"q4f16_0/{fname}"
        out[key + "/y"] = y
        out[key + "/meta"] = np.array([K, N, seed, qo.EPILOGUES[ep]], dtype=np.int64)
        out[key + "/sha"] = np.frombuffer(sha(packed, scale, x, epv).encode(), dtype=np.uint8)
        out[key + "/secs"] = np.array([dt])
        print(f"fused {fname}: {dt:.2f}s  |y|max={np.abs(y.astype(np.float64)).max():.3f}")

    # int3 "before" kernels (dispatch/llama/main.py:5316-5758); import replays 77 schedules (~12 s)
    try:
        from mlc_llm.dispatch.llama import main as lm

        fmt3 = qo.FORMATS["q3f16_0"]
        names = [n for n in dir(lm) if n.endswith("int3_fp16_before")]
        print("int3 before funcs:", names)
        for ci, fname in enumerate(sorted(names)):
            pf = getattr(lm, fname)
            bufs = [pf.buffer_map[p] for p in pf.params]
            shapes = [tuple(int(s) for s in b.shape) for b in bufs]
            dts = [b.dtype for b in bufs]
            R, N = shapes[0]
            K = shapes[2][-1]
            seed = 7100 + ci
            packed, scale, x, epv = synth_fused_inputs(fmt3, K, N, seed)
            assert packed.shape == (R, N), (packed.shape, shapes)
            args = [packed, scale, x.reshape(shapes[2])]
            if len(shapes) == 5:
                args.append(epv.reshape(shapes[3]))
            args.append(np.zeros(shapes[-1], dts[-1]))
            res, dt = build_run(tvm, pf, args)
            key = f"q3f16_0/{fname}"
            out[key + "/y"] = res[-1].reshape(-1)
            out[key + "/meta"] = np.array([K, N, seed, len(shapes)], dtype=np.int64)
            out[key + "/sha"] = np.frombuffer(sha(packed, scale, x, epv).encode(), dtype=np.uint8)
            print(f"fused {fname}: K={K} N={N} {dt:.2f}s")
    except Exception as e:  # pragma: no cover
        print("int3 before funcs skipped:", repr(e))
    np.savez_compressed(os.path.join(GOLDEN, "fused_gemv.npz"), **out)


def gen_asym(tvm, refq):
    """The reference CLI's DEFAULT scheme (asymmetric int4; quantization.py:105-140 encode, :183-211 decode,
    :235-255 take_decode) and its fused GEMVs (dispatch/llama/main.py:4397-4415 f32, :4979-4997 f16)."""
    from tvm import te

    out = {}
    for ci, (name, N, K) in enumerate([("q4f16_asym", 24, 256), ("q4f16_asym", 16, 4096), ("q4f32_asym", 24, 256),
                                       ("q4f32_asym", 8, 2560)]):
        fmt = qo.ASYM_FORMATS[name]
        f32 = fmt.dtype == "float32"
        rng = np.random.default_rng(2000 + ci)
        W = (rng.standard_normal((N, K), dtype=np.float32) * 0.02).astype(fmt.np_dtype)
        W[1, 0] = fmt.np_dtype(0.5)
        W[2, 1] = fmt.np_dtype(-0.25)
        R, G = K // 8, K // 32
        aux_dt = np.uint32 if f32 else fmt.np_dtype
        res = {"W": W}
        for tr in (True, False):
            enc = refq.encoding_func(False, 32, 4, "int4", 32, transpose=tr, dtype=fmt.dtype)
            Wt = te.placeholder((N, K), fmt.dtype, name="W")
            outs = list(enc(Wt))
            f = te.create_prim_func([Wt] + outs)
            pshape, ashape = ((R, N), (G, N)) if tr else ((N, R), (N, G))
            bufs = [W, np.zeros(pshape, np.uint32)] + [np.zeros(ashape, aux_dt) for _ in outs[1:]]
            r, _ = build_run(tvm, f, bufs)
            sfx = "" if tr else "_rm"
            res["packed" + sfx] = r[1]
            res["scale" + sfx] = r[2]
            if not f32:
                res["min" + sfx] = r[3]
        aux = [res["scale"]] + ([] if f32 else [res["min"]])
        for tr in (False, True):
            dec = refq.decoding_func(False, 32, 4, "int4", 32, K, data_transposed=True, transpose_output=tr, dtype=fmt.dtype)
            D = te.placeholder((R, N), "uint32", name="D")
            A = [te.placeholder((G, N), "uint32" if f32 else fmt.dtype, name=f"A{i}") for i in range(len(aux))]
            f = te.create_prim_func([D] + A + [dec(D, *A)])
            r, _ = build_run(tvm, f, [res["packed"]] + aux + [np.zeros((N, K) if tr else (K, N), fmt.np_dtype)])
            res["w_t" if tr else "w"] = r[-1]
        # arbitrary words / aux values
        rw = rng.integers(0, 2**32, size=(R, N), dtype=np.uint64).astype(np.uint32)
        if f32:
            raux = [qo.f32x2_to_bf16x2((rng.random((G, N), dtype=np.float32) * 0.01), (-rng.random((G, N), dtype=np.float32) * 0.08))]
        else:
            raux = [(rng.random((G, N), dtype=np.float32) * 0.01).astype(np.float16), (-rng.random((G, N), dtype=np.float32) * 0.08).astype(np.float16)]
        dec = refq.decoding_func(False, 32, 4, "int4", 32, K, data_transposed=True, transpose_output=False, dtype=fmt.dtype)
        D = te.placeholder((R, N), "uint32", name="D")
        A = [te.placeholder((G, N), "uint32" if f32 else fmt.dtype, name=f"A{i}") for i in range(len(raux))]
        f = te.create_prim_func([D] + A + [dec(D, *A)])
        r, _ = build_run(tvm, f, [rw] + raux + [np.zeros((K, N), fmt.np_dtype)])
        res.update(rand_packed=rw, rand_scale=raux[0], rand_w=r[-1])
        if not f32:
            res["rand_min"] = raux[1]
        # embedding gather
        ids = np.array([3, 0, N - 1, 3, 7], dtype=np.int32)
        tk = refq.decoding_after_taking_func(False, 32, 4, "int4", 32, K, dtype=fmt.dtype)
        D = te.placeholder((N, R), "uint32", name="D")
        A = [te.placeholder((N, G), "uint32" if f32 else fmt.dtype, name=f"A{i}") for i in range(len(aux))]
        I = te.placeholder((len(ids),), "int32", name="I")
        f = te.create_prim_func([D] + A + [I, tk(D, *A, I)])
        aux_rm = [res["scale_rm"]] + ([] if f32 else [res["min_rm"]])
        r, _ = build_run(tvm, f, [res["packed_rm"]] + aux_rm + [ids, np.zeros((len(ids), K), fmt.np_dtype)])
        res.update(take_ids=ids, take_out=r[-1])
        for k_, v_ in res.items():
            out[f"{name}_N{N}_K{K}/{k_}"] = v_
        print(f"asym encode/decode {name} N={N} K={K}: ok")
    # fused GEMVs, the reference's unscheduled PrimFuncs at 4096 x 4096
    try:
        sys.path.insert(0, REF)
        import mlc_llm  # noqa: F401
        from mlc_llm.dispatch.llama import main as lm

        for ci, (fname, name) in enumerate([("fused_decode4_matmul5_before", "q4f32_asym"), ("fused_decode4_matmul5_fp16_before", "q4f16_asym")]):
            fmt = qo.ASYM_FORMATS[name]
            f32 = fmt.dtype == "float32"
            K = N = 4096
            rng = np.random.default_rng(7300 + ci)
            packed = rng.integers(0, 2**32, size=(K // 8, N), dtype=np.uint64).astype(np.uint32)
            sc = rng.random((K // 32, N), dtype=np.float32) * 0.01
            mn = -rng.random((K // 32, N), dtype=np.float32) * 0.08
            aux = [qo.f32x2_to_bf16x2(sc, mn)] if f32 else [sc.astype(np.float16), mn.astype(np.float16)]
            x = rng.standard_normal(K, dtype=np.float32).astype(fmt.np_dtype)
            res, dt = build_run(tvm, getattr(lm, fname), [packed] + aux + [x.reshape(1, 1, K), np.zeros((1, 1, N), fmt.np_dtype)])
            key = f"{name}/{fname}"
            out[key + "/y"] = res[-1].reshape(N)
            out[key + "/seed"] = np.array([7300 + ci], dtype=np.int64)
            out[key + "/sha"] = np.frombuffer(sha(packed, *aux, x).encode(), dtype=np.uint8)
            print(f"asym fused {fname}: {dt:.2f}s")
    except Exception as e:  # pragma: no cover
        print("asym fused skipped:", repr(e))
    np.savez_compressed(os.path.join(GOLDEN, "asym.npz"), **out)


def gen_variants(tvm, refq):
    """uint16-storage int3 and fp4 (quantization.py:51-101, 142-177, 213-270 with storage_nbit=16 / mode="fp4")."""
    from tvm import te

    out = {}
    for ci, (name, N, K) in enumerate([("q3f16_u16", 24, 320), ("q3f16_u16", 8, 4096), ("q3f16_u16", 24, 100),
                                       ("fp4f16_0", 24, 256), ("fp4f32_0", 24, 256), ("fp4f16_0", 8, 4096)]):
        fmt = qo.VARIANT_FORMATS[name]
        mode = "fp4" if qo.is_fp4(fmt) else "int3"
        sdt = "uint16" if fmt.storage_nbit == 16 else "uint32"
        rng = np.random.default_rng(3000 + ci)
        W = (rng.standard_normal((N, K), dtype=np.float32) * 0.02).astype(fmt.np_dtype)
        W[0, : fmt.group] = 0
        W[1, 0] = fmt.np_dtype(0.5)
        R, G = fmt.n_words(K), fmt.n_groups(K)
        res = {"W": W}
        for tr in (True, False):
            enc = refq.encoding_func(True, fmt.group, fmt.nbit, mode, fmt.storage_nbit, transpose=tr, dtype=fmt.dtype)
            Wt = te.placeholder((N, K), fmt.dtype, name="W")
            pt, st = enc(Wt)
            f = te.create_prim_func([Wt, pt, st])
            pshape, sshape = ((R, N), (G, N)) if tr else ((N, R), (N, G))
            (_, p_, s_), _ = build_run(tvm, f, [W, np.zeros(pshape, sdt), np.zeros(sshape, fmt.np_dtype)])
            res["packed" + ("" if tr else "_rm")] = p_
            res["scale" + ("" if tr else "_rm")] = s_
        if name == "fp4f16_0":
            # the reference's own fp4 -> f16 decoder (_tir_u32_to_f4_to_f16, quantization.py:88-101) reinterprets a
            # uint32 expression as float16; TVM rejects the size mismatch, so that PrimFunc cannot be built: only the
            # encoder is pinned for this scheme (the decode restatement takes the evident intent, the low 16 bits)
            for k_, v_ in res.items():
                out[f"{name}_N{N}_K{K}/{k_}"] = v_
            print(f"variant {name} N={N} K={K}: encode only (reference decoder does not build)")
            continue
        for tr in (False, True):
            dec = refq.decoding_func(True, fmt.group, fmt.nbit, mode, fmt.storage_nbit, K, data_transposed=True,
                                     transpose_output=tr, dtype=fmt.dtype)
            D = te.placeholder((R, N), sdt, name="D")
            S = te.placeholder((G, N), fmt.dtype, name="S")
            f = te.create_prim_func([D, S, dec(D, S)])
            (_, _, w), _ = build_run(tvm, f, [res["packed"], res["scale"], np.zeros((N, K) if tr else (K, N), fmt.np_dtype)])
            res["w_t" if tr else "w"] = w
        rw = rng.integers(0, 2**fmt.storage_nbit, size=(R, N), dtype=np.uint64).astype(sdt)
        rs = (rng.random((G, N), dtype=np.float32) * 0.01).astype(fmt.np_dtype)
        dec = refq.decoding_func(True, fmt.group, fmt.nbit, mode, fmt.storage_nbit, K, data_transposed=True,
                                 transpose_output=False, dtype=fmt.dtype)
        D = te.placeholder((R, N), sdt, name="D")
        S = te.placeholder((G, N), fmt.dtype, name="S")
        f = te.create_prim_func([D, S, dec(D, S)])
        (_, _, w), _ = build_run(tvm, f, [rw, rs, np.zeros((K, N), fmt.np_dtype)])
        res.update(rand_packed=rw, rand_scale=rs, rand_w=w)
        if K % fmt.group == 0:
            ids = np.array([3, 0, N - 1, 3, 7], dtype=np.int32)
            tk = refq.decoding_after_taking_func(True, fmt.group, fmt.nbit, mode, fmt.storage_nbit, K, dtype=fmt.dtype)
            D = te.placeholder((N, R), sdt, name="D")
            S = te.placeholder((N, G), fmt.dtype, name="S")
            I = te.placeholder((len(ids),), "int32", name="I")
            f = te.create_prim_func([D, S, I, tk(D, S, I)])
            (_, _, _, emb), _ = build_run(tvm, f, [res["packed_rm"], res["scale_rm"], ids, np.zeros((len(ids), K), fmt.np_dtype)])
            res.update(take_ids=ids, take_out=emb)
        for k_, v_ in res.items():
            out[f"{name}_N{N}_K{K}/{k_}"] = v_
        print(f"variant {name} N={N} K={K}: ok")
    np.savez_compressed(os.path.join(GOLDEN, "variants.npz"), **out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--skip-fused", action="store_true")
    ap.add_argument("--skip-encdec", action="store_true")
    ap.add_argument("--only-asym", action="store_true")
    ap.add_argument("--only-variants", action="store_true")
    a = ap.parse_args()
    os.makedirs(GOLDEN, exist_ok=True)
    tvm = _load_tvm()
    refq = _load_ref_quant()
    if a.only_asym:
        gen_asym(tvm, refq)
        return
    if a.only_variants:
        gen_variants(tvm, refq)
        return
    gen_asym(tvm, refq)
    gen_variants(tvm, refq)
    if not a.skip_encdec:
        gen_encode_decode(tvm, refq)
    if not a.skip_fused:
        gen_fused(tvm)


if __name__ == "__main__":
    main()
