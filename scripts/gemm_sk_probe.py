"""stream-K experiments on two shapes: python scripts/gemm_sk_probe.py  (B200Q_GEMM_DEBUG / B200Q_GEMM_STREAMK in the environment)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS
dev = torch.device("cuda", 0)
M = int(os.environ.get("M", 2048))
def t(fn, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
spec = SPECS["q4f16_0"]
for K, N in ((2560, 2560), (10240, 2560), (2560, 10240), (4096, 4096), (2560, 7680)):
    p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
    sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
    X = (torch.randn(M, K, device=dev) * 0.05).half()
    Y = torch.empty(M, N, dtype=torch.float16, device=dev)
    fl = 2.0 * M * K * N
    a = t(lambda: ops.gemm(p, sc, X, "q4f16_0", out=Y))
    b = t(lambda: ops.gemm(p, sc, X, "q4f16_0", out=Y, ws=False))
    print(f"SK dbg={os.environ.get('B200Q_GEMM_DEBUG', '0')} {M}x{K}x{N}: stream-K {a * 1e3:7.1f} us {fl / a / 1e9:7.1f} TF | whole {b * 1e3:7.1f} us {fl / b / 1e9:7.1f} TF", flush=True)
