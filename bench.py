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

N > 1 (torchrun): tensor parallel - q/k/v/gate/up column-sharded, o/down row-sharded, one all-reduce
per sharded block (total work fixed => "scaling": "strong").

Prints ONE JSON line (rank 0).
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
# CPU baseline / reference arm: the reference's algorithm (f16-accumulating fused GEMV) on host cores
# ------------------------------------------------------------------------------------------------
def cpu_sample(model: str, fmt_name: str):
    """One layer of the workload (its 7 quantized Linears), weights packed by the oracle encoder."""
    import numpy as np

    from mlc_llm_b200.decode_stack import MODEL_SHAPES, linear_list, token_macs
    from oracle import quant_oracle as qo

    shape = MODEL_SHAPES[model]
    fmt = qo.FORMATS[fmt_name]
    ops_ = []
    for j, (name, K, N, ep) in enumerate(linear_list(shape)):
        W = qo.synth_weight(N, K, fmt, seed=qo.SEED + j)
        packed, scale = qo.encode_sym(W, fmt)
        x = qo.synth_act(K, fmt, seed=qo.SEED + 100 + j)
        ep0 = qo.synth_act(N, fmt, seed=qo.SEED + 200 + j)
        ops_.append((packed, scale, x, qo.EPILOGUES[ep], ep0))
    macs = sum(K * N for _, K, N, _ in linear_list(shape))
    return fmt, ops_, macs / token_macs(shape)


def cpu_run_sample(fmt, ops_):
    from oracle import c_oracle as co

    t = time.perf_counter()
    for packed, scale, x, ep, ep0 in ops_:
        co.gemv_ref(packed, scale, x, fmt, ep, ep0)
    return time.perf_counter() - t


def cpu_baseline(model: str, fmt_name: str, steps: int = 3, warmup: int = 1):
    from oracle import c_oracle as co

    fmt, ops_, frac = cpu_sample(model, fmt_name)
    for _ in range(warmup):
        cpu_run_sample(fmt, ops_)
    ts = [cpu_run_sample(fmt, ops_) for _ in range(steps)]
    t = sum(ts) / len(ts)
    return {
        "value": frac / t,
        "unit": UNIT,
        "cores": co.num_threads(),
        "kind": "port",
        "sample": f"one decoder layer (7 fused dequant-GEMVs, {frac * 100:.2f}% of a token's MACs), "
                  f"{steps} runs, {t:.2f} s each, scaled to a whole token; oracle/oracle.c (C/OpenMP port of the "
                  "reference's f16-accumulating fused_decode*_matmul* loop nest)",
    }, t


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps = max(1, min(args.steps, 5))
    warmup = max(1, min(args.warmup, 1))
    cb, t = cpu_baseline(args.model, args.fmt, steps=steps, warmup=warmup)
    out = {
        "impl": "reference",
        "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": steps, "warmup": warmup,
        "ms_per_step": 1000.0 / cb["value"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f16", "data": "synthetic", "config": workload_config(args, 1),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference arm = the reference's CPU algorithm (build.py --target llvm hands LLVM the unscheduled loop "
                "nest; LLVM-enabled TVM is not available, so the C/OpenMP port in oracle/ is timed on all host cores); "
                "ms_per_step is per whole token, extrapolated from the bounded sample",
    }
    print(json.dumps(out))


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

    stack = DecodeStack(shape, spec, device=dev, tp_rank=rank, tp_size=world, layers=args.layers,
                        pdl=not args.no_pdl, fuse_qkv=not args.no_fuse_qkv, fuse_gate_up=not args.no_fuse_gate_up,
                        use_program=(args.fmt == "q4f16_0") if args.program < 0 else bool(args.program))
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
    t_wall1 = time.time()
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
        sampler.stop()
        clocks = sampler.summary(t_wall0, t_wall2)

    # ---- per-shape GB/s (weights rotate through all layers => every byte from HBM) -----------------
    per_shape = []
    if rank == 0 and world == 1 and not args.no_per_shape and stack.n_layers >= 8:
        per_shape = bench_shapes(stack, spec, torch)

    # ---- roofline of the dominant kernel (the GEMV) -----------------------------------------------
    step_bytes = token_bytes(shape, spec, world) if args.layers is None else None
    peak, peak_src = measured_peaks()
    launches = stack.launches_per_step
    roofline = None
    if step_bytes is not None:
        achieved = step_bytes / (ms_per_step * 1e-3) / 1e9
        roofline = {
            "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": ncu_traffic("gemv_program_tc_llama7b_token_dram_bytes") if (stack.use_program and args.fmt == "q4f16_0" and world == 1 and args.model == "llama-7b") else None,
            "kernel": (("b200q::tc::gemv_program_tc_kernel" if args.fmt == "q4f16_0" else "b200q::gemv_program_kernel") +
                       " (persistent: every GEMV of the step in one cooperative launch per collective-free segment" +
                       ("; tcgen05.mma with the unpacked weights as TMEM operand" if args.fmt == "q4f16_0" else "") + ")") if stack.use_program else "b200q::gemv_stream_kernel (one launch per fused Linear group, PDL chain)",
            "gemv_steps_per_launch": (sum(o.n_steps for k, o in stack._plan if k == "program") / max(1, launches)) if stack.use_program else 1,
            "bytes_per_step_per_rank": step_bytes, "launches_per_step": launches,
            "avg_launch_us": ms_per_step * 1e3 / launches,
            "peak_source": peak_src + (" of measured" if "measured" in peak_src else " of fallback"),
            "note": "achieved = algorithmic bytes of the step's GEMV launches (packed + scales + x + out + epilogue operand, "
                    "SURVEY.md 8d) / CUDA-event time of the step" + (("; the per-block reductions run inside the program launch (peer pushes over NVLink), the logits all-gather is NCCL" if getattr(stack, "fused_tp", False) else "; includes the all-reduces") if world > 1 else ""),
        }

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
            "gpu_launches": int(ops.launch_count() - n0) if stack.graph is None else launches * args.steps,
            "collectives_per_step": stack.collectives_per_step,
            "clocks": clocks,
            "aggregate_gbs": (step_bytes * world / (ms_per_step * 1e-3) / 1e9) if step_bytes else None,
            "per_shape": per_shape,
        }
        if args.layers is not None:
            out["invalid"] = "debug run with --layers: not the named config"
        if world == 1 and not args.no_cpu_baseline:
            cb, _ = cpu_baseline(args.model, args.fmt, steps=2, warmup=1)
            out["cpu_baseline"] = cb
        print(json.dumps(out))
    if world > 1:
        # leave without tearing NCCL down: destroy_process_group() can block forever while CUDA graphs that
        # captured collectives are still alive; every rank has finished its work at this barrier
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
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
