"""A few launches of the prefill dequant-GEMM pair kernel for ncu: q3f16_0 M = 2048, (K, N) = (10240, 2560) and
(2560, 2560), whole-tile schedule then stream-K (ncu: -k regex:gemm_tc2 -s 8 -c 4 captures one of each)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS
dev = torch.device("cuda", 0)
M, fmt = 2048, "q3f16_0"
spec = SPECS[fmt]
cases = []
for K, N in ((10240, 2560), (2560, 2560)):
    p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
    sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
    X = (torch.randn(M, K, device=dev) * 0.05).half()
    Y = torch.empty(M, N, dtype=torch.float16, device=dev)
    cases.append((p, sc, X, Y))
ops.gemm_workspace(dev)
for rep in range(3):  # 2 warm rounds (8 launches), then the captured one
    for p, sc, X, Y in cases:
        ops.gemm(p, sc, X, fmt, out=Y, ws=False)   # whole tiles
        ops.gemm(p, sc, X, fmt, out=Y)             # stream-K (chosen by the cost model for both shapes)
torch.cuda.synchronize()
print("done")
