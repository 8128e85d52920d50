"""prefill Linears of config 3 (dolly-v2-3b q3f16_0, M = 2048) with their epilogues, against the same GEMM without one"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS
dev = torch.device("cuda", 0)
M, fmt = 2048, "q3f16_0"
spec = SPECS[fmt]
def t(fn, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for K, N, ep in ((2560, 2560, "bias"), (2560, 10240, "bias_gelu"), (10240, 2560, "bias")):
    p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
    sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
    X = (torch.randn(M, K, device=dev) * 0.05).half()
    b = (torch.randn(N, device=dev) * 0.1).half()
    Y = torch.empty(M, N, dtype=torch.float16, device=dev)
    fl = 2.0 * M * K * N
    a = t(lambda: ops.gemm(p, sc, X, fmt, out=Y))
    c = t(lambda: ops.gemm(p, sc, X, fmt, ep, b, out=Y))
    d = t(lambda: ops.gemm(p, sc, X, fmt, ep, b, out=Y, pdl=True))
    e = t(lambda: ops.gemm(p, sc, X, fmt, ep, b, out=Y, pdl=True, streamk=False))
    print(f"EP {fmt} {M}x{K}x{N}: none {a * 1e3:6.1f} us {fl / a / 1e9:6.1f} TF | {ep} {c * 1e3:6.1f} us {fl / c / 1e9:6.1f} TF | +pdl {d * 1e3:6.1f} us {fl / d / 1e9:6.1f} TF | +pdl whole tiles {e * 1e3:6.1f} us {fl / e / 1e9:6.1f} TF", flush=True)
