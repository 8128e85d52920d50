#!/usr/bin/env python
"""Experiment driver: prefill GEMM throughput per shape (CUDA events, L2 flushed between iterations
is unnecessary here: the kernel is tensor-bound and X + packed W are L2-sized by design)."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--fmts", default="q3f16_0,q4f16_0,q8f16_0")
    ap.add_argument("--shapes", default="2048x2560x2560,2048x2560x10240,2048x10240x2560,2048x4096x4096,2048x4096x11008,2048x11008x4096,8192x8192x8192")
    ap.add_argument("--cublas", type=int, default=1)
    args = ap.parse_args()
    dev = torch.device("cuda")
    for sh in args.shapes.split(","):
        M, K, N = [int(v) for v in sh.split("x")]
        X = (torch.randn(M, K, device=dev)).half()
        fl = 2.0 * M * K * N
        if args.cublas:
            T = (torch.randn(N, K, device=dev) * 0.02).half()
            us = timeit(lambda: torch.matmul(X, T.t()))
            print(f"{sh:20s} cuBLAS f16 dense   : {us:9.1f} us {fl / us / 1e6:8.1f} TFLOP/s", flush=True)
            us = timeit(lambda: ops.nt_matmul(X, T))
            print(f"{sh:20s} b200q nt_matmul    : {us:9.1f} us {fl / us / 1e6:8.1f} TFLOP/s", flush=True)
        for name in args.fmts.split(","):
            spec = SPECS[name]
            p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
            s = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
            out = torch.empty((M, N), dtype=torch.float16, device=dev)
            us = timeit(lambda: ops.gemm(p, s, X, name, out=out))
            print(f"{sh:20s} b200q gemm {name:8s}: {us:9.1f} us {fl / us / 1e6:8.1f} TFLOP/s", flush=True)
            if args.cublas:
                Td = torch.empty((N, K), dtype=torch.float16, device=dev)
                us2 = timeit(lambda: (ops.dequantize(p, s, K, name, True, out=Td), ops.nt_matmul(X, Td, out=out)))
                print(f"{sh:20s} 2-kernel   {name:8s}: {us2:9.1f} us {fl / us2 / 1e6:8.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
