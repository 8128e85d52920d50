"""time of the stream-K / whole-tile pair kernel against the work per pair (M = 2048, K = 2560, N swept)"""
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
spec = SPECS["q4f16_0"]
M = 2048
for K in (2560, 5120):
    for N in (256, 1280, 2304, 2560, 3840, 4864, 5120, 7680):
        p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
        sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
        X = (torch.randn(M, K, device=dev) * 0.05).half()
        Y = torch.empty(M, N, dtype=torch.float16, device=dev)
        tiles = 8 * ((N + 255) // 256)
        a = t(lambda: ops.gemm(p, sc, X, "q4f16_0", out=Y))
        b = t(lambda: ops.gemm(p, sc, X, "q4f16_0", out=Y, ws=False))
        print(f"SWEEP K={K} N={N}: tiles {tiles} ({tiles / 74:.2f}/pair) stream-K {a * 1e3:6.1f} us  whole {b * 1e3:6.1f} us", flush=True)
