#!/usr/bin/env python
"""Experiment driver (not part of the product): per-shape GEMV bandwidth under b200q_gemv_tune
settings, weights rotating through a pool larger than L2, CUDA-graph replay, PDL on."""
import argparse
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS


def pool(spec, K, N, n_mats, dev):
    R, G = spec.n_words(K), spec.n_groups(K)
    ws = []
    for i in range(n_mats):
        p = torch.randint(-2**31, 2**31 - 1, (R, N), dtype=torch.int32, device=dev)
        s = (torch.rand((G, N), device=dev) * 0.01).to(spec.torch_dtype)
        ws.append((p, s))
    return ws


def time_graph(fn, iters=20):
    st = torch.cuda.Stream()
    st.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(st):
        fn(); st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            fn()
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
    return e0.elapsed_time(e1) * 1e3 / iters  # us per graph


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fmt", default="q4f16_0")
    ap.add_argument("--shapes", default="4096x4096,4096x11008,11008x4096,4096x22016,4096x32000")
    ap.add_argument("--tunes", default="0,0,0;0,0,1;2,0,0;2,0,1;0,3,0;0,8,0")
    ap.add_argument("--pdl", type=int, default=1)
    ap.add_argument("--chain", type=int, default=0, help="1: every op reads the previous op's output (K==N only)")
    args = ap.parse_args()
    dev = torch.device("cuda")
    spec = SPECS[args.fmt]
    ws = ops.workspace()
    for sh in args.shapes.split(","):
        K, N = [int(v) for v in sh.split("x")]
        nbytes = spec.gemv_bytes(K, N)
        n_mats = max(8, int(600e6 // nbytes) + 1)
        mats = pool(spec, K, N, n_mats, dev)
        x = (torch.randn(K, device=dev) * 0.05).to(spec.torch_dtype)
        outs = [torch.empty(N, dtype=spec.torch_dtype, device=dev) for _ in range(2)]
        for tune in args.tunes.split(";"):
            t = [int(v) for v in tune.split(",")]
            ops.gemv_tune(*t)
            def fn():
                for i, (p, s) in enumerate(mats):
                    xin = x
                    if args.chain and K == N and i > 0:
                        xin = outs[(i - 1) & 1]
                    ops.gemv(p, s, xin, spec, out=outs[i & 1], pdl=bool(args.pdl), ws=ws)
            us = time_graph(fn) / n_mats
            print(f"{args.fmt} K={K} N={N} tune={tune:8s} pdl={args.pdl} chain={args.chain}: {us:8.2f} us  {nbytes / us / 1e3:8.1f} GB/s  ({nbytes/1e6:.1f} MB x {n_mats})", flush=True)
        ops.gemv_tune(0, 0, 0)
        del mats


if __name__ == "__main__":
    main()
