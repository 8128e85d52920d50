#!/usr/bin/env python
"""Experiment driver: single-step decode programs on big matrices (consumer compute rate of the
persistent kernel) - weights rotate through a pool larger than L2 inside ONE program."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS

dev = torch.device("cuda")
fmt = sys.argv[1] if len(sys.argv) > 1 else "q4f16_0"
spec = SPECS[fmt]
for sh in (sys.argv[2] if len(sys.argv) > 2 else "4096x4096,4096x12288,4096x22016,11008x4096,4096x32000").split(","):
    K, N = [int(v) for v in sh.split("x")]
    nbytes = spec.gemv_bytes(K, N)
    n = max(8, int(600e6 // nbytes) + 1)
    x = (torch.randn(K, device=dev) * 0.05).to(spec.torch_dtype)
    steps = []
    for i in range(n):
        p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
        s = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).to(spec.torch_dtype)
        steps.append(dict(weights=[(p, s)], x=x, outs=[torch.empty(N, dtype=spec.torch_dtype, device=dev)]))
    prog = ops.GemvProgram(steps, fmt)   # independent steps: no dependency bubbles
    for _ in range(3):
        prog.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        prog.run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 10 / n
    print(f"{fmt} K={K} N={N}: {us:8.2f} us/step {nbytes / us / 1e3:8.1f} GB/s ({n} independent steps in one launch)", flush=True)
