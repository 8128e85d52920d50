"""TFLOP/s of the prefill dequant-GEMM per tile kernel (single-CTA 128x256, CTA-pair 256x128 / 256x256, auto) and of
cuBLAS on the dequantized weight, M = 2048."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS

dev = torch.device("cuda", 0)
M = 2048
def t(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for fmt in ("q3f16_0", "q4f16_0", "q8f16_0"):
    spec = SPECS[fmt]
    for K, N in ((2560, 2560), (2560, 10240), (10240, 2560), (4096, 11008), (4096, 4096)):
        if fmt == "q3f16_0" and K % 40: continue
        p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
        sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
        X = (torch.randn(M, K, device=dev) * 0.05).half()
        Y = torch.empty(M, N, dtype=torch.float16, device=dev)
        fl = 2.0 * M * K * N
        res = []
        for tile in ("single", "pair128", "pair256", "whole", None):
            try:
                if tile == "whole":  # the pair kernel's whole-tile schedule (no stream-K)
                    ms = t(lambda: ops.gemm(p, sc, X, fmt, out=Y, ws=False))
                else:
                    ms = t(lambda: ops.gemm(p, sc, X, fmt, out=Y, tile=tile, ws=(None if tile is None else False)))
                res.append(f"{tile or 'auto(stream-K)'} {fl / ms / 1e9:7.1f}")
            except Exception as e:
                res.append(f"{tile} ERR {str(e)[:60]}")
        Wt = ops.dequantize(p, sc, K, fmt, True)
        ms = t(lambda: torch.matmul(X, Wt.t(), out=Y))
        res.append(f"cublas-dense {fl / ms / 1e9:7.1f}")
        ms = t(lambda: ops.nt_matmul(X, Wt, out=Y))
        res.append(f"nt_matmul {fl / ms / 1e9:7.1f}")
        print(f"GEMM {fmt} {M}x{K}x{N}: " + " | ".join(res), flush=True)
