"""q | k | v of config 3 (q3f16_0, 2048 x 2560 -> 3 x 2560, +bias): three launches against one b200q_gemm_multi launch"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS
dev = torch.device("cuda", 0)
def t(fn, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for fmt, M, K, N in (("q3f16_0", 2048, 2560, 2560), ("q4f16_0", 2048, 4096, 4096), ("q4f16_0", 2048, 8192, 1024)):
    spec = SPECS[fmt]
    W = []
    for i in range(3):
        p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
        sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
        W.append((p, sc))
    X = (torch.randn(M, K, device=dev) * 0.05).half()
    b = [(torch.randn(N, device=dev) * 0.1).half() for _ in range(3)]
    Y = [torch.empty(M, N, dtype=torch.float16, device=dev) for _ in range(3)]
    fl = 3 * 2.0 * M * K * N
    def three():
        for i in range(3):
            ops.gemm(W[i][0], W[i][1], X, fmt, "bias", b[i], out=Y[i])
    a = t(three)
    c = t(lambda: ops.gemm_multi(W, X, fmt, ["bias"] * 3, b, outs=Y))
    print(f"MULTI {fmt} {M}x{K}x3x{N}: three launches {a * 1e3:6.1f} us {fl / a / 1e9:6.1f} TF | one launch {c * 1e3:6.1f} us {fl / c / 1e9:6.1f} TF", flush=True)
