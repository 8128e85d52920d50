"""One launch each of the prefill dequant-GEMM kernels (for ncu): q4f16_0 2048 x 4096 x 11008, tiles single / pair256,
then q3f16_0 2048 x 2560 x 10240 pair256 and the dense pair256."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.quant_spec import SPECS
dev = torch.device("cuda", 0)
M = 2048
for fmt, K, N, tiles in (("q4f16_0", 4096, 11008, ("single", "pair256")), ("q3f16_0", 2560, 10240, ("pair256",))):
    spec = SPECS[fmt]
    p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
    sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
    X = (torch.randn(M, K, device=dev) * 0.05).half()
    Y = torch.empty(M, N, dtype=torch.float16, device=dev)
    for tile in tiles:
        for _ in range(2):
            ops.gemm(p, sc, X, fmt, out=Y, tile=tile)
    if fmt == "q4f16_0":
        Wt = ops.dequantize(p, sc, K, fmt, True)
        for _ in range(2):
            ops.nt_matmul(X, Wt, out=Y)
torch.cuda.synchronize()
print("done")
