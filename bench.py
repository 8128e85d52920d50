#!/usr/bin/env python
"""bench.py - the decode-step quantized-Linear stream of BASELINE.json's configs[1]
(vicuna-v1-7b = Llama-7B shape, q4f16_0, batch 1) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path (oracle port)

One "step" = one decode token = every quantized-Linear launch of the model in the reference's order
(SURVEY.md 3.2): 32 x (q|k|v, o+residual, gate+silu, up*gate, down+residual) + lm_head.  Attention,
norms and sampling are not part of this path and are NOT in the timed region (stated in `config`).
The 3.7 GB of distinct packed weights is far larger than the 126 MB L2, so every step streams its
weights from HBM (config.l2: "inputs larger than L2").

N > 1 (torchrun): tensor parallel - q/k/v/gate/up column-sharded, o/down row-sharded, one reduction
per sharded block (total work fixed => "scaling": "strong").

Prints ONE JSON line (rank 0).  Besides the contract's keys the line carries, all measured OUTSIDE the timed
region of the headline:
  parity     a 2-layer stack of the same shapes on the same ranks through the same kernel path, every fused
             dequant-GEMV re-evaluated by the CPU oracle from the bits it read (oracle/stack_check.py)
  segmented  the same token with the program cut at the reference's non-GEMV call sites (what a drop-in behind
             the reference's operator boundary gets today)
  prefill    BASELINE config 3: dolly-v2-3b q3f16_0, 2048 prompt tokens, tcgen05 dequant-GEMM, TFLOP/s
  extra      BASELINE configs 1 / 4 / 5 as they fit the N of this run
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "llama7b_q4f16_0_batch1_decode_linear_stream_tokens_per_s"
UNIT = "tokens/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="llama-7b")
    ap.add_argument("--fmt", default="q4f16_0")
    ap.add_argument("--layers", type=int, default=None, help="debug: fewer layers (result is marked invalid)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-per-shape", action="store_true")
    ap.add_argument("--no-pdl", action="store_true")
    ap.add_argument("--program", type=int, default=-1, help="1: the persistent decode-program kernel (one cooperative launch per collective-free segment; tcgen05 kernel for q4f16_0); 0: one launch per fused Linear group, PDL chain; -1: auto (program for q4f16_0)")
    ap.add_argument("--no-fuse-qkv", action="store_true")
    ap.add_argument("--no-fuse-gate-up", action="store_true")
    ap.add_argument("--tune", default="", help="ctas_per_sm,stages,dbg_flags for b200q_gemv_tune (experiments)")
    ap.add_argument("--segments", default="fused", choices=["fused", "reference"],
                    help="headline launch granularity: fused = one program per token (default); reference = cut at the "
                         "reference's non-GEMV call sites (the other one is reported beside it either way)")
    ap.add_argument("--headline-only", action="store_true", help="skip parity / segmented / prefill / extra records")
    ap.add_argument("--no-prefill", action="store_true")
    ap.add_argument("--no-extra", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--ref-budget-s", type=float, default=150.0, help="wall-clock budget of the whole --impl reference run")
    return ap.parse_args()


def workload_config(args, n_gpus):
    return {
        "workload": f"{args.model} {args.fmt} batch-1 decode: quantized-Linear stream of one token "
                    "(225 reference GEMVs; attention/norm/sampling excluded: out of the path's scope)",
        "model_shape": args.model,
        "format": args.fmt,
        "batch": 1,
        "parallelism": "single" if n_gpus == 1 else f"tp{n_gpus} (column q/k/v/gate/up, row o/down, all-reduce per block)",
        "l2": "inputs larger than L2 (distinct packed weights per step >> 126 MB)",
        "weights": "random-init N(0,0.02^2), packed by the library encoder (bit-exact vs the reference encoder)",
    }


# ------------------------------------------------------------------------------------------------
# clocks sampling (B200_PROFILING.md): nvidia-smi during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.samples, self.proc, self.thread = index, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()

    def summary(self, t0: float, t1: float):
        sm, mx, reasons, pw = [], 0.0, set(), []
        for t, line in self.samples:
            if t < t0 or t > t1 + 0.15:
                continue
            f = [v.strip() for v in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "power_w_max": max(pw) if pw else None}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic(key: str):
    """dram bytes per launch of the dominant kernel from the committed ncu --set full capture."""
    p = os.path.join(ROOT, "profiles", "roofline_traffic.json")
    if os.path.exists(p):
        return json.load(open(p)).get(key)
    return None


# ------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's algorithm (dtype-accumulating fused GEMV) on host cores
# ------------------------------------------------------------------------------------------------
def host_threads() -> int:
    """Host cores this process may use (torchrun exports OMP_NUM_THREADS=1: never trust the OpenMP default)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


class CpuSample:
    """A bounded sample of one token's quantized-Linear work for the CPU port: EVERY Linear of one decoder layer
    plus 1/L of lm_head's output columns = exactly 1/L of a token (L layers), weights packed by the oracle
    encoder.  `col_frac` < 1 evaluates only the leading fraction of every operator's output columns (the cost of
    the reference's loop nest is proportional to N), so that a run of many steps stays inside its time budget."""

    def __init__(self, model: str, fmt_name: str):
        import numpy as np

        from mlc_llm_b200.decode_stack import MODEL_SHAPES, linear_list, token_macs
        from oracle import quant_oracle as qo

        self.shape = MODEL_SHAPES[model]
        self.fmt = qo.FORMATS[fmt_name]
        self.token_macs = token_macs(self.shape)
        items = list(linear_list(self.shape))
        n_lm = max(8, (self.shape.vocab // self.shape.layers) // 8 * 8)
        items.append(("lm_head", self.shape.hidden, n_lm, "none" if self.fmt.dtype == "float32" else "cast_f32"))
        self.ops = []
        for j, (name, K, N, ep) in enumerate(items):
            W = qo.synth_weight(N, K, self.fmt, seed=qo.SEED + j)
            packed, scale = qo.encode_sym(W, self.fmt)
            x = qo.synth_act(K, self.fmt, seed=qo.SEED + 100 + j)
            ep0 = qo.synth_act(N, self.fmt, seed=qo.SEED + 200 + j)
            self.ops.append((name, K, N, packed, scale, x, qo.EPILOGUES[ep], ep0))
        self.col_frac = 1.0
        self._np = np
        self._sliced = None

    def set_col_frac(self, f: float):
        self.col_frac = min(1.0, max(1.0 / 64, f))
        self._sliced = None

    def _prepare(self):
        np = self._np
        out = []
        macs = 0
        for name, K, N, packed, scale, x, ep, ep0 in self.ops:
            n = N if self.col_frac >= 1.0 else max(8, int(N * self.col_frac) // 8 * 8)
            out.append((np.ascontiguousarray(packed[:, :n]), np.ascontiguousarray(scale[:, :n]), x, ep, np.ascontiguousarray(ep0[:n])))
            macs += K * n
        self._sliced = out
        self.fraction_of_token = macs / self.token_macs

    def run(self) -> float:
        from oracle import c_oracle as co

        if self._sliced is None:
            self._prepare()
        t = time.perf_counter()
        for packed, scale, x, ep, ep0 in self._sliced:
            co.gemv_ref(packed, scale, x, self.fmt, ep, ep0)
        return time.perf_counter() - t

    def describe(self, steps: int, t: float) -> str:
        cf = "" if self.col_frac >= 1.0 else f", leading {self.col_frac * 100:.1f}% of every operator's output columns"
        return (f"every quantized Linear of one decoder layer + 1/{self.shape.layers} of lm_head{cf} = "
                f"{self.fraction_of_token * 100:.3f}% of a token's MACs per step, {steps} timed steps, {t:.3f} s each, "
                "tokens/s = fraction / time; oracle/oracle.c (C/OpenMP port of the reference's dtype-accumulating "
                "fused_decode*_matmul* loop nest, mlc_llm/dispatch/dispatch_tir_operator_adreno.py:5-23)")


def cpu_baseline(model: str, fmt_name: str, steps: int = 3, warmup: int = 1, budget_s: float = 20.0):
    from oracle import c_oracle as co

    co.set_threads(host_threads())
    sample = CpuSample(model, fmt_name)
    t1 = sample.run()  # calibration (also the first warm-up)
    if (steps + warmup) * t1 > budget_s:
        sample.set_col_frac(budget_s / ((steps + warmup) * t1))
    for _ in range(max(0, warmup - 1)):
        sample.run()
    ts = [sample.run() for _ in range(steps)]
    t = sum(ts) / len(ts)
    return {
        "value": sample.fraction_of_token / t,
        "unit": UNIT,
        "cores": co.num_threads(),
        "kind": "port",
        "sample": sample.describe(steps, t),
    }, t, sample


def run_reference(args):
    """The reference arm: the reference's own CPU algorithm for the path, all host threads, K steps after W
    warm-up steps, each step a bounded sample of the workload (see CpuSample)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    t_begin = time.time()
    cb, t, sample = cpu_baseline(args.model, args.fmt, steps=steps, warmup=max(1, warmup), budget_s=args.ref_budget_s)
    out = {
        "impl": "reference",
        "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f16" if sample.fmt.dtype == "float16" else "f32", "data": "synthetic",
        "config": workload_config(args, args.gpus),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "sample_fraction_of_token": sample.fraction_of_token,
        "wall_s": round(time.time() - t_begin, 2),
        "note": "reference arm = the reference's CPU algorithm (build.py --target llvm hands LLVM the unscheduled loop "
                "nest; an LLVM-enabled TVM is not available here, so the C/OpenMP port in oracle/ is timed on all host "
                "threads - it is FASTER than the reference's own TVM-C build of the same PrimFunc, which runs one "
                "4096x4096 GEMV in ~2 s on one thread); one step = the bounded sample described in cpu_baseline.sample "
                "(ms_per_step is the measured time of that sample; value = sample_fraction_of_token / step time); the host "
                "path does not change with --gpus: rank 0 runs it with every host thread at every N",
    }
    print(json.dumps(out))


# ------------------------------------------------------------------------------------------------
# records measured outside the headline's timed region
# ------------------------------------------------------------------------------------------------
def time_steps(torch, fn, steps: int, warmup: int, barrier, max_over_ranks) -> float:
    """ms per call of fn(): CUDA events on the current stream, barrier + synchronize on both sides, max over ranks."""
    for _ in range(max(warmup, 3)):
        fn()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    barrier()
    return max_over_ranks(e0.elapsed_time(e1)) / steps


def parity_record(args, torch, dist, dev, rank, world, use_program: bool):
    """A 2-layer stack of the SAME shapes on the SAME ranks through the SAME kernel path as the headline; every
    fused dequant-GEMV it launched is re-evaluated by the CPU oracle (oracle/stack_check.py) from the bits it
    read.  At N > 1 the shards are gathered and checked against the full (unsharded) weights regenerated on
    rank 0 from the same seeds - the only way a tensor-parallel line can carry correctness."""
    from mlc_llm_b200 import ops
    from mlc_llm_b200.decode_stack import MODEL_SHAPES, DecodeStack

    shape = MODEL_SHAPES[args.model]
    st = DecodeStack(shape, args.fmt, device=dev, tp_rank=rank, tp_size=world, layers=2, use_program=use_program,
                     pdl=not args.no_pdl, segments=args.segments)
    g = torch.Generator(device=dev)
    g.manual_seed(1234)
    st.x_in.copy_((torch.randn(shape.hidden, device=dev, generator=g) * 0.05).to(st.dt))
    st.capture()
    st.decode_step()
    st.decode_step()
    snap = st.snapshot()
    group_scaled = bool(use_program and ops.program_uses_tensor_cores(args.fmt))
    rec = None
    if rank == 0:
        from oracle import quant_oracle as qo
        from oracle import stack_check

        full = st if world == 1 else DecodeStack(shape, args.fmt, device=dev, layers=2, use_program=False)

        def weights(li, nm):
            w = full.lm_head if nm == "lm_head" else full.layers[li][nm]
            return w.packed.cpu().numpy().view("uint32"), w.scale.cpu().numpy()

        rec = stack_check.check_launches(snap, weights, qo.FORMATS[args.fmt], group_scaled)
        rec["what"] = (f"{args.model} shapes, 2 layers + lm_head, {args.fmt}, " + ("program" if use_program else "PDL chain") +
                       (f", tp{world} (fused in-program reduction)" if getattr(st, "fused_tp", False) else (f", tp{world} (NCCL all-reduce)" if world > 1 else "")) +
                       ": every launched fused dequant-GEMV vs the f64 oracle on the bits it read")
        del full
    del st
    torch.cuda.empty_cache()
    return rec


def prefill_record(args, torch, dev, clocks_fn):
    """BASELINE config 3: dolly-v2-3b (GPT-NeoX shape), q3f16_0, 2048 prompt tokens: the 6 Linears x 32 layers of one
    `encoding` call through the tcgen05 dequant-GEMM (ops.gemm -> b200q_gemm) with their fused epilogues; beside
    it the reference's own CUDA route rebuilt from library parts (b200q_dequant to [N,K] + cuBLAS NT matmul +
    the same epilogue as torch ops) and the CPU port on a row sample."""
    from mlc_llm_b200 import ops
    from mlc_llm_b200.decode_stack import MODEL_SHAPES, DecodeStack
    from mlc_llm_b200.prefill_stack import PrefillStack

    model, fmt, M = "dolly-v2-3b", "q3f16_0", 2048
    st = DecodeStack(MODEL_SHAPES[model], fmt, device=dev, use_program=False)
    pf = PrefillStack(st, M)
    g = torch.Generator(device=dev)
    g.manual_seed(7)
    pf.x_in.copy_((torch.randn(M, st.shape.hidden, device=dev, generator=g) * 0.05).half())
    ident = lambda v: v
    t_wall0 = time.time()
    n0 = ops.launch_count()
    ms = time_steps(torch, pf.run, 10, 3, torch.cuda.synchronize, ident)
    launches = (ops.launch_count() - n0) // 13
    clocks = clocks_fn(t_wall0, time.time())
    flops = pf.flops()
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("bf16_tflops_sustained", 1379.4))
    burst = float(peaks.get("bf16_tflops", 1639.4))
    tf = flops / (ms * 1e-3) / 1e12
    rec = {
        "metric": "dolly_v2_3b_q3f16_0_prefill2048_linear_tflops", "value": tf, "unit": "TFLOP/s", "ms_per_pass": ms,
        "config": {"workload": f"{model} {fmt} prefill, M={M}: 6 quantized Linears x {st.n_layers} layers with fused epilogues "
                               "(bias, bias+GELU, bias+2 residual adds); attention / norms excluded", "flop": flops,
                   "l2": "192 Linears (q|k|v|h_to_4h of a layer in one launch: 96 launches) over 1.1 GB of distinct packed weights per pass"},
        "roofline": {"bound": "tensor", "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak,
                     "frac_of_burst": tf / burst, "kernel": "b200q::gemm_tc2_kernel<q3f16_0, 256> (CTA pair, tcgen05.mma.cta_group::2 M = 256, dequant in the B-operand warps, stream-K schedule when whole tiles fill the pairs badly)",
                     "flop_per_launch": flops / max(1, launches), "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (burst beside it)"},
        "gpu_launches_per_pass": launches, "clocks": clocks,
    }
    # per-shape TFLOP/s of the three distinct GEMMs
    shapes = []
    w0 = st.layers[0]
    for label, w, ep in (("2560->2560 +bias", w0["q"], "bias"), ("2560->10240 +bias+gelu", w0["h_to_4h"], "bias_gelu"),
                         ("10240->2560 +bias", w0["4h_to_h"], "bias")):
        X = (torch.randn(M, w.K, device=dev, generator=g) * 0.05).half()
        bias = torch.zeros(w.N, dtype=torch.float16, device=dev)
        Y = torch.empty(M, w.N, dtype=torch.float16, device=dev)
        f = lambda: ops.gemm(w.packed, w.scale, X, fmt, ep, bias, out=Y, pdl=True)  # (as PrefillStack launches it)
        t = time_steps(torch, f, 20, 3, torch.cuda.synchronize, ident)
        # the reference's CUDA route from library parts: dequantize to [N, K] in HBM, then a dense NT matmul (cuBLAS)
        Wt = torch.empty(w.N, w.K, dtype=torch.float16, device=dev)

        def ref_route():
            ops.dequantize(w.packed, w.scale, w.K, fmt, True, out=Wt)
            torch.matmul(X, Wt.t(), out=Y)

        t_ref = time_steps(torch, ref_route, 20, 3, torch.cuda.synchronize, ident)
        fl = 2.0 * M * w.K * w.N
        shapes.append({"shape": f"{M}x{label}", "us": round(t * 1e3, 2), "tflops": round(fl / t / 1e9, 1),
                       "dequant_plus_cublas_us": round(t_ref * 1e3, 2), "dequant_plus_cublas_tflops": round(fl / t_ref / 1e9, 1)})
    # q | k | v of a layer as the stack runs them: ONE launch (b200q_gemm_multi) against three, and the plain GEMM
    # (no epilogue) of the widest Linear - the GELU costs the epilogue warps ~25 us of a 2560 -> 10240 Linear
    X = (torch.randn(M, w0["q"].K, device=dev, generator=g) * 0.05).half()
    Ys = [torch.empty(M, w0[n].N, dtype=torch.float16, device=dev) for n in "qkv"]
    bs = [torch.zeros(w0[n].N, dtype=torch.float16, device=dev) for n in "qkv"]
    Ws = [(w0[n].packed, w0[n].scale) for n in "qkv"]
    t_one = time_steps(torch, lambda: ops.gemm_multi(Ws, X, fmt, ["bias"] * 3, bs, outs=Ys, pdl=True), 20, 3, torch.cuda.synchronize, ident)

    def three():
        for i in range(3):
            ops.gemm(Ws[i][0], Ws[i][1], X, fmt, "bias", bs[i], out=Ys[i], pdl=True)

    t_three = time_steps(torch, three, 20, 3, torch.cuda.synchronize, ident)
    fl = 2.0 * M * w0["q"].K * sum(w0[n].N for n in "qkv")
    shapes.append({"shape": f"{M}x2560->3x2560 +bias (q|k|v, one launch)", "us": round(t_one * 1e3, 2), "tflops": round(fl / t_one / 1e9, 1),
                   "three_launches_us": round(t_three * 1e3, 2), "three_launches_tflops": round(fl / t_three / 1e9, 1)})
    w = w0["h_to_4h"]
    X = (torch.randn(M, w.K, device=dev, generator=g) * 0.05).half()
    Y = torch.empty(M, w.N, dtype=torch.float16, device=dev)
    t = time_steps(torch, lambda: ops.gemm(w.packed, w.scale, X, fmt, out=Y, pdl=True), 20, 3, torch.cuda.synchronize, ident)
    shapes.append({"shape": f"{M}x2560->10240 (no epilogue)", "us": round(t * 1e3, 2), "tflops": round(2.0 * M * w.K * w.N / t / 1e9, 1)})
    rec["per_shape"] = shapes
    # CPU port on a row sample: the reference's f16-accumulating NT matmul loop nest on the dequantized weight
    if not args.no_cpu_baseline:
        import numpy as np

        from oracle import c_oracle as co
        from oracle import quant_oracle as qo

        co.set_threads(host_threads())
        f3 = qo.FORMATS[fmt]
        rows = 8
        tot, macs = 0.0, 0
        for w in (w0["q"], w0["h_to_4h"], w0["4h_to_h"]):
            pk, sc = w.packed.cpu().numpy().view(np.uint32), w.scale.cpu().numpy()
            T = co.decode(pk, sc, w.K, f3, transpose_output=True)
            Xs = qo.synth_act((rows, w.K), f3, seed=5)
            co.nt_matmul_ref_f16(Xs, T)
            t0 = time.perf_counter()
            co.nt_matmul_ref_f16(Xs, T)
            tot += time.perf_counter() - t0
            macs += rows * w.K * w.N
        rec["cpu_baseline"] = {"value": 2.0 * macs / tot / 1e12, "unit": "TFLOP/s", "cores": co.num_threads(), "kind": "port",
                               "sample": f"{rows} of the {M} rows through the three distinct GEMM shapes of one layer (NT_matmul only, weight "
                                         "already dequantized), oracle/oracle.c f16-accumulating loop nest"}
    del pf, st
    torch.cuda.empty_cache()
    return rec


def reference_gpu_record(torch, dev):
    """The reference's OWN hand-scheduled kernels (TVM-generated CUDA of mlc_llm/dispatch PrimFuncs, built for sm_100a
    by oracle/build_ref_cuda.py) on this GPU, beside this library's kernel for the same operator, same protocol:
    32 distinct weight sets per sweep (>= 300 MB: every byte from HBM), CUDA events, back-to-back launches."""
    from mlc_llm_b200 import ops
    from mlc_llm_b200.quant_spec import SPECS
    from oracle import ref_cuda

    if not ref_cuda.available():
        return {"unavailable": "oracle/_ref/cuda not built (oracle/build_ref_cuda.py needs /root/reference + the TVM bundled with tilelang)"}
    man = ref_cuda.manifest()
    ident = lambda v: v
    rows = []
    R = 32
    h = lambda *s: (torch.randn(*s, device=dev) * 0.05).half()
    rnd_w = lambda r, n: torch.randint(-2**31, 2**31 - 1, (r, n), dtype=torch.int32, device=dev)
    rnd_s = lambda g, n: (torch.rand((g, n), device=dev) * 0.01).half()
    # (kernel, K, N, our spec, epilogue, which operands the reference kernel takes)
    gemvs = [("gemv_q4f16_0_4096x4096", 4096, 4096, "q4f16_0", "none", False),
             ("gemv_q4f16_0_4096x11008_silu", 4096, 11008, "q4f16_0", "silu", False),
             ("gemv_q4f16_asym_4096x4096", 4096, 4096, "q4f16_asym", "none", True)]
    for name, K, N, fmt, ep, has_min in gemvs:
        if name not in man:
            continue
        spec = SPECS[fmt]
        k = ref_cuda.RefKernel(name)
        ws = [(rnd_w(K // 8, N), rnd_s(K // 32, N), rnd_s(K // 32, N) if has_min else None) for _ in range(R)]
        x, out = h(1, 1, K), torch.zeros((1, 1, N), dtype=torch.float16, device=dev)

        def ref_sweep():
            for p, sc, mn in ws:
                k(*([p, sc] + ([mn] if has_min else []) + [x, out]))

        def our_sweep():
            for p, sc, mn in ws:
                ops.gemv(p, sc, x, fmt, ep, out=out, pdl=True, zmin=mn)

        t_ref = time_steps(torch, ref_sweep, 5, 3, torch.cuda.synchronize, ident) * 1e3 / R
        t_our = time_steps(torch, our_sweep, 5, 3, torch.cuda.synchronize, ident) * 1e3 / R
        nb = spec.gemv_bytes(K, N, ep)
        rows.append({"op": f"{fmt} GEMV {K}->{N} {ep}", "reference_kernel": name, "grid": man[name]["grid"], "block": man[name]["block"],
                     "reference_us": round(t_ref, 2), "reference_gbs": round(nb / t_ref / 1e3, 1), "b200q_us": round(t_our, 2),
                     "b200q_gbs": round(nb / t_our / 1e3, 1), "speedup": round(t_ref / t_our, 2)})
    if "decode_q4f16_asym_4096x4096" in man and "nt_matmul_f16_4096x4096" in man:
        K = N = 4096
        M = 2048
        kd, km = ref_cuda.RefKernel("decode_q4f16_asym_4096x4096"), ref_cuda.RefKernel("nt_matmul_f16_4096x4096")
        p, sc, mn = rnd_w(K // 8, N), rnd_s(K // 32, N), rnd_s(K // 32, N)
        X, Wt = h(1, M, K), torch.zeros((N, K), dtype=torch.float16, device=dev)
        Y = torch.zeros((1, M, N), dtype=torch.float16, device=dev)

        def ref_prefill():  # the reference's CUDA route: decode to [N, K] in HBM, then the SIMT NT_matmul
            kd(p, sc, mn, Wt)
            km(X, Wt, Y)

        t_ref = time_steps(torch, ref_prefill, 5, 3, torch.cuda.synchronize, ident) * 1e3
        t_our = time_steps(torch, lambda: ops.gemm(p, sc, X[0], "q4f16_asym", out=Y[0], zmin=mn), 10, 3, torch.cuda.synchronize, ident) * 1e3
        fl = 2.0 * M * K * N
        rows.append({"op": f"q4f16_asym prefill Linear {M}x{K}->{N} (decode4_fp16 + NT_matmul1_fp16 vs one tcgen05 dequant-GEMM)",
                     "reference_us": round(t_ref, 1), "reference_tflops": round(fl / t_ref / 1e6, 2), "b200q_us": round(t_our, 1),
                     "b200q_tflops": round(fl / t_our / 1e6, 1), "speedup": round(t_ref / t_our, 1)})
    return {"what": "the reference's hand-scheduled PrimFuncs lowered to CUDA by TVM 0.23 and compiled with nvcc for sm_100a "
                    "(oracle/build_ref_cuda.py); launch geometry is the schedule's own", "rows": rows}


def decode_extra(args, torch, dist, dev, rank, world, model, fmt, barrier, max_over_ranks, steps=30, cpu=False):
    """One more decode configuration of BASELINE.json on the ranks of this run (same protocol, fewer steps)."""
    from mlc_llm_b200 import ops
    from mlc_llm_b200.decode_stack import MODEL_SHAPES, DecodeStack, token_bytes
    from mlc_llm_b200.quant_spec import SPECS

    shape, spec = MODEL_SHAPES[model], SPECS[fmt]
    # the one-launch program where it runs on the tensor cores (q4f16_0, q8f16_0); the per-operator PDL chain otherwise
    # (q4f32_0: 696 against 277 tok/s through the register-level program kernel)
    st = DecodeStack(shape, spec, device=dev, tp_rank=rank, tp_size=world, use_program=ops.program_uses_tensor_cores(spec))
    st.x_in.copy_((torch.randn(shape.hidden, device=dev) * 0.05).to(st.dt))
    st.capture()
    ms = time_steps(torch, st.decode_step, steps, 5, barrier, max_over_ranks)
    nbytes = token_bytes(shape, spec, world)
    peak, _ = measured_peaks()
    rec = {"metric": f"{model}_{fmt}_batch1_decode_linear_stream_tokens_per_s", "value": 1000.0 / ms, "unit": UNIT,
           "n_gpus": world, "ms_per_step": ms, "steps": steps,
           "config": {"workload": f"{model} {fmt} batch-1 decode: quantized-Linear stream of one token", "parallelism": "single" if world == 1 else f"tp{world}",
                      "reduction": ("in-program (peer pushes over NVLink)" if getattr(st, "fused_tp", False) else ("NCCL all-reduce per sharded block" if world > 1 else "none"))},
           "bytes_per_step_per_rank": nbytes, "gbs_per_rank": nbytes / ms / 1e6, "frac_of_hbm_peak": nbytes / ms / 1e6 / peak,
           "launches_per_step": st.launches_per_step, "collectives_per_step": st.collectives_per_step}
    if cpu and rank == 0 and not args.no_cpu_baseline:
        cb, _, _ = cpu_baseline(model, fmt, steps=2, warmup=1, budget_s=15.0)
        rec["cpu_baseline"] = cb
    del st
    torch.cuda.empty_cache()
    return rec


def prefill_extra(args, torch, dist, dev, rank, world, model, fmt, barrier, max_over_ranks, M=2048, steps=3):
    """Tensor-parallel prefill (BASELINE config 5): column-sharded q/k/v/gate/up GEMMs, row-sharded o/down GEMMs whose
    per-rank partial sums are summed by one NCCL all-reduce per sharded block."""
    from mlc_llm_b200.decode_stack import MODEL_SHAPES, DecodeStack
    from mlc_llm_b200.prefill_stack import PrefillStack

    st = DecodeStack(MODEL_SHAPES[model], fmt, device=dev, tp_rank=rank, tp_size=world, use_program=False)
    pf = PrefillStack(st, M)
    pf.x_in.copy_((torch.randn(M, st.shape.hidden, device=dev) * 0.05).half())
    ms = time_steps(torch, pf.run, steps, 3, barrier, max_over_ranks)
    flops = pf.flops() * world
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("bf16_tflops_sustained", 1379.4)) * world
    rec = {"metric": f"{model}_{fmt}_prefill{M}_linear_tflops", "value": flops / ms / 1e9, "unit": "TFLOP/s (aggregate)", "n_gpus": world,
           "ms_per_pass": ms, "config": {"workload": f"{model} {fmt} prefill M={M}, tp{world}: sharded tcgen05 dequant-GEMMs + one NCCL all-reduce of "
                                                     f"[{M}, {st.shape.hidden}] f16 per sharded block", "flop": flops},
           "frac_of_tensor_peak": flops / ms / 1e9 / peak, "gemms_per_pass": pf.gemms_per_pass, "collectives_per_pass": pf.collectives_per_pass}
    del pf, st
    torch.cuda.empty_cache()
    return rec


# ------------------------------------------------------------------------------------------------
def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
        return

    import torch
    import torch.distributed as dist

    from mlc_llm_b200 import ops
    from mlc_llm_b200.decode_stack import MODEL_SHAPES, DecodeStack, linear_list, token_bytes
    from mlc_llm_b200.quant_spec import SPECS

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n_gpus = world
    spec = SPECS[args.fmt]
    shape = MODEL_SHAPES[args.model]
    use_program = ops.program_uses_tensor_cores(args.fmt) if args.program < 0 else bool(args.program)

    stack = DecodeStack(shape, spec, device=dev, tp_rank=rank, tp_size=world, layers=args.layers,
                        pdl=not args.no_pdl, fuse_qkv=not args.no_fuse_qkv, fuse_gate_up=not args.no_fuse_gate_up,
                        use_program=use_program, segments=args.segments)
    if args.tune:
        ops.gemv_tune(*[int(v) for v in args.tune.split(",")])
    stack.x_in.copy_((torch.randn(shape.hidden, device=dev) * 0.05).to(stack.dt))
    stack.capture()
    torch.cuda.synchronize()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms: float) -> float:
        if world == 1:
            return ms
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)

    # ---- device-resident timing -------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        stack.decode_step()
    barrier()
    n0 = ops.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.time()
    ev0.record()
    for _ in range(args.steps):
        stack.decode_step()
    ev1.record()
    barrier()
    ms = max_over_ranks(ev0.elapsed_time(ev1))
    ms_per_step = ms / args.steps
    tok_s = 1000.0 / ms_per_step

    # ---- end to end through the public API, host buffers ---------------------------------------------
    x_host = (torch.randn(shape.hidden) * 0.05).to(stack.dt).pin_memory()
    logits_host = torch.empty(shape.vocab, dtype=torch.float32).pin_memory()
    for _ in range(3):
        stack.decode_step_host(x_host, logits_host)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        stack.decode_step_host(x_host, logits_host)
    e1.record()
    barrier()
    e2e_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    t_wall2 = time.time()

    clocks = None
    if rank == 0:
        time.sleep(0.15)
        clocks = sampler.summary(t_wall0, t_wall2)

    def clocks_between(t0, t1):
        time.sleep(0.15)
        return sampler.summary(t0, t1) if rank == 0 else None

    full_run = args.layers is None and not args.headline_only
    launches = stack.launches_per_step
    collectives = stack.collectives_per_step

    # ---- the other launch granularity beside the headline (same weights, same protocol) ----------------
    segmented = None
    if full_run and use_program and world == 1:
        other = "reference" if args.segments == "fused" else "fused"
        st2 = DecodeStack(shape, spec, device=dev, use_program=True, segments=other, weights_from=stack)
        st2.x_in.copy_(stack.x_in)
        st2.capture()
        ms2 = time_steps(torch, st2.decode_step, min(args.steps, 100), 5, barrier, max_over_ranks)
        segmented = {"segments": other, "value": 1000.0 / ms2, "unit": UNIT, "ms_per_step": ms2,
                     "launches_per_step": st2.launches_per_step,
                     "what": ("program cut wherever the reference's `decoding` has a non-GEMV call_tir between two fused GEMVs "
                              "(after q|k|v, after o, after down: mlc_llm/relax_model/llama.py:340-401): 3 launches per layer + lm_head, "
                              "every segment starts from a plain buffer - what a drop-in behind the reference's operator boundary "
                              "gets today") if other == "reference" else "one program launch per token"}
        if (stack.logits - st2.logits).abs().max().item() != 0.0:
            segmented["logits_equal_to_headline"] = False
        else:
            segmented["logits_equal_to_headline"] = True
        del st2
        torch.cuda.empty_cache()

    # ---- per-shape GB/s (weights rotate through all layers => every byte from HBM) -----------------
    per_shape = []
    if rank == 0 and world == 1 and not args.no_per_shape and stack.n_layers >= 8 and not args.headline_only:
        per_shape = bench_shapes(stack, spec, torch)

    # ---- roofline of the dominant kernel (the GEMV) -----------------------------------------------
    step_bytes = token_bytes(shape, spec, world) if args.layers is None else None
    peak, peak_src = measured_peaks()
    roofline = None
    if step_bytes is not None:
        achieved = step_bytes / (ms_per_step * 1e-3) / 1e9
        tc = use_program and ops.program_uses_tensor_cores(args.fmt)
        roofline = {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": ncu_traffic("gemv_program_tc_llama7b_token_dram_bytes") if (tc and world == 1 and args.model == "llama-7b" and args.segments == "fused") else None,
            "traffic_source": "profiles/roofline_traffic.json (dram__bytes_read + dram__bytes_write of one launch, ncu --set full of this command; not re-measured by this run)",
            "kernel": (("b200q::tc::gemv_program_tc_kernel" if tc else "b200q::gemv_program_kernel") +
                       " (persistent: every GEMV of the step in one cooperative launch per collective-free segment" +
                       ("; tcgen05.mma with the unpacked weights as TMEM operand" if tc else "") + ")") if use_program else "b200q::gemv_stream_kernel (one launch per fused Linear group, PDL chain)",
            "gemv_steps_per_launch": (sum(o.n_steps for k, o in stack._plan if k == "program") / max(1, launches)) if use_program else 1,
            "bytes_per_step_per_rank": step_bytes, "launches_per_step": launches,
            "avg_launch_us": ms_per_step * 1e3 / max(1, launches),
            "peak_source": peak_src + (" of measured" if "measured" in peak_src else " of fallback"),
            "note": "achieved = algorithmic bytes of the step's GEMV launches (packed + scales + x + out + epilogue operand, "
                    "SURVEY.md 8d) / CUDA-event time of the step" + (("; the per-block reductions run inside the program launch (peer pushes over NVLink), the logits all-gather is NCCL" if getattr(stack, "fused_tp", False) else "; includes the all-reduces") if world > 1 else ""),
        }
    graph_used = stack.graph is not None
    n_launched = int(ops.launch_count() - n0)

    # ---- records outside the timed region ------------------------------------------------------------------
    del stack
    torch.cuda.empty_cache()
    parity = None
    if not args.no_parity and not args.headline_only:
        try:
            parity = parity_record(args, torch, dist, dev, rank, world, use_program)
        except Exception as e:  # noqa: BLE001
            parity = {"ok": False, "error": f"{type(e).__name__}: {str(e)[:300]}"}
    prefill = None
    if full_run and world == 1 and not args.no_prefill:
        try:
            prefill = prefill_record(args, torch, dev, clocks_between)
        except Exception as e:  # noqa: BLE001
            prefill = {"error": f"{type(e).__name__}: {str(e)[:300]}"}
    reference_gpu = None
    if full_run and world == 1:
        try:
            reference_gpu = reference_gpu_record(torch, dev)
        except Exception as e:  # noqa: BLE001
            reference_gpu = {"error": f"{type(e).__name__}: {str(e)[:300]}"}
    extra = []

    def guarded(fn, *a, **kw):
        # a failure in one of the side records must never cost the headline line (every rank takes the same path: the
        # records are collective-free except where every rank calls them in the same order)
        try:
            extra.append(fn(*a, **kw))
        except Exception as e:  # noqa: BLE001
            extra.append({"error": f"{fn.__name__}{a[7:9]}: {type(e).__name__}: {str(e)[:300]}"})
            torch.cuda.empty_cache()

    if full_run and not args.no_extra:
        if world == 1 and (args.model, args.fmt) != ("redpajama-3b", "q4f32_0"):
            guarded(decode_extra, args, torch, dist, dev, rank, world, "redpajama-3b", "q4f32_0", barrier, max_over_ranks, cpu=True)
        if (args.model, args.fmt) != ("llama-13b", "q8f16_0"):
            guarded(decode_extra, args, torch, dist, dev, rank, world, "llama-13b", "q8f16_0", barrier, max_over_ranks)
        if world == 8:
            guarded(decode_extra, args, torch, dist, dev, rank, world, "llama-65b", "q4f16_0", barrier, max_over_ranks, steps=20)
            guarded(prefill_extra, args, torch, dist, dev, rank, world, "llama-65b", "q4f16_0", barrier, max_over_ranks)
    if rank == 0:
        sampler.stop()

    if rank == 0:
        out = {
            "metric": METRIC, "value": tok_s, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f16 (f32 accumulate)", "data": "synthetic",
            "config": workload_config(args, n_gpus),
            "roofline": roofline,
            "e2e": {"value": 1000.0 / e2e_ms, "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": x_host.numel() * x_host.element_size(),
                    "d2h_bytes_per_step": logits_host.numel() * logits_host.element_size(),
                    "api": "DecodeStack.decode_step_host(x_host, logits_host): pinned hidden state in, f32 logits out, stream sync"},
            "gpu_launches": n_launched if not graph_used else launches * args.steps,
            "collectives_per_step": collectives,
            "clocks": clocks,
            "aggregate_gbs": (step_bytes * world / (ms_per_step * 1e-3) / 1e9) if step_bytes else None,
            "parity": parity,
            "segmented": segmented,
            "per_shape": per_shape,
            "prefill": prefill,
            "reference_gpu": reference_gpu,
            "extra": extra,
        }
        if args.layers is not None:
            out["invalid"] = "debug run with --layers: not the named config"
        if world == 1 and not args.no_cpu_baseline:
            cb, _, _ = cpu_baseline(args.model, args.fmt, steps=3, warmup=1, budget_s=20.0)
            out["cpu_baseline"] = cb
        print(json.dumps(out))
    if world > 1:
        # every CUDA graph that captured a collective (and every symmetric-memory workspace) was released above;
        # tear NCCL down cleanly, but never let a stuck teardown hang the run: every rank has finished its work at
        # this barrier and the JSON line is out
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        done = threading.Event()

        def _teardown():
            try:
                dist.destroy_process_group()
            finally:
                done.set()

        threading.Thread(target=_teardown, daemon=True).start()
        if not done.wait(timeout=20.0):
            os._exit(0)


def bench_shapes(stack, spec, torch, iters: int = 20):
    """Per-shape GEMV time with the weights rotating through every layer (>= 300 MB per sweep)."""
    from mlc_llm_b200 import ops

    res = []
    L = stack.n_layers
    cases = {}
    if stack.shape.family != "llama":
        return res
    h, i = stack.shape.hidden, stack.shape.inter
    cases[f"q|k|v 3x({h}->{h}) one launch"] = ("multi", ["q", "k", "v"], "x_in", None)
    cases[f"o {h}->{h} +residual"] = ("single", "o", "q", ("add", "x1"))
    cases[f"gate {h}->{i} silu"] = ("single", "gate", "x1", ("silu", None))
    cases[f"up {h}->{i} *gate"] = ("single", "up", "x1", ("multiply", "g"))
    cases[f"gate|up 2x({h}->{i}) one launch"] = ("multi", ["gate", "up"], "x1", None)
    cases[f"down {i}->{h} +residual"] = ("single", "down", "a", ("add", "x1"))
    for label, (kind, names, xname, ep) in cases.items():
        def sweep():
            for li in range(L):
                w, b = stack.layers[li], stack.buf[li]
                x = stack.x_in if xname == "x_in" else b[xname]
                if kind == "multi":
                    ops.gemv_multi([(w[n].packed, w[n].scale) for n in names], x, spec,
                                   outs=[b["g"], b["u"]] if names[0] == "gate" else [b[n] for n in names], pdl=True,
                                   ws=stack.ws)
                else:
                    ops.gemv(w[names].packed, w[names].scale, x, spec, ep[0], b[ep[1]] if ep[1] else None,
                             out=b["x2"] if w[names].N == stack.shape.hidden else b["a"], pdl=True, ws=stack.ws)
        st = torch.cuda.Stream()
        st.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(st):
            sweep()
            st.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=st):
                sweep()
        torch.cuda.current_stream().wait_stream(st)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (iters * L)
        if kind == "multi":
            nbytes = sum(spec.gemv_bytes(stack.layers[0][n].K, stack.layers[0][n].N, "none") for n in names)
        else:
            nbytes = spec.gemv_bytes(stack.layers[0][names].K, stack.layers[0][names].N, ep[0])
        res.append({"shape": label, "us": round(us, 3), "gbs": round(nbytes / us / 1e3, 1), "bytes": nbytes})
    # lm_head: single 73.8 MB matrix (< L2): rotate by interleaving with a >L2 sweep is not possible with one
    # matrix, so it is timed inside the whole step only; reported here from an L2-flushed single launch
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=stack.device)
    ts = []
    x = stack.buf[-1]["x2"]
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.gemv(stack.lm_head.packed, stack.lm_head.scale, x, spec, "cast_f32", None, out=stack.logits_part, ws=stack.ws)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    nbytes = spec.gemv_bytes(stack.lm_head.K, stack.lm_head.N, "cast_f32")
    us = ts[len(ts) // 2]
    res.append({"shape": f"lm_head {stack.lm_head.K}->{stack.lm_head.N} f32 (L2 flushed, event-timed single launch)", "us": round(us, 3),
                "gbs": round(nbytes / us / 1e3, 1), "bytes": nbytes})
    if spec.name == "q4f16_0":
        res.append(stream_rate(spec, torch, stack.device))
    return res


def stream_rate(spec, torch, dev, K: int = 4096, N: int = 32000, n: int = 9, iters: int = 10):
    """Streaming rate of the tcgen05 decode-program kernel without dependency bubbles: n INDEPENDENT
    K x N GEMVs (distinct weights, n * 73.8 MB >> L2) as the steps of one program launch."""
    from mlc_llm_b200 import ops
    x = (torch.randn(K, device=dev) * 0.05).to(spec.torch_dtype)
    steps = []
    for _ in range(n):
        p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
        sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).to(spec.torch_dtype)
        steps.append(dict(weights=[(p, sc)], x=x, outs=[torch.empty(N, dtype=spec.torch_dtype, device=dev)]))
    prog = ops.GemvProgram(steps, spec.name)
    for _ in range(3):
        prog.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        prog.run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters / n
    nbytes = spec.gemv_bytes(K, N)
    peak, _ = measured_peaks()
    return {"shape": f"stream: {n} independent {K}->{N} GEMVs as steps of one program launch (no dependency bubbles, weights >> L2)",
            "us": round(us, 3), "gbs": round(nbytes / us / 1e3, 1), "bytes": nbytes, "frac_of_peak": round(nbytes / us / 1e3 / peak, 4)}


if __name__ == "__main__":
    main()
