import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from mlc_llm_b200 import ops
from util import dev, host
K, N = 128, 128
q = np.full((K, N), 7)
for k in range(K):
    for n in range(N):
        if k % 32 == n % 32 and k // 32 == n // 32: q[k, n] = 8
R = K // 8
packed = np.zeros((R, N), dtype=np.uint32)
for t in range(8):
    packed |= (q[t::8, :].astype(np.uint32) & 15) << (4 * t)
scale = np.ones((K // 32, N), dtype=np.float16)
pd, sd = dev(packed), dev(scale)
for k0 in list(range(0, 34)) + [64, 127]:
    x = np.zeros(K, dtype=np.float16); x[k0] = 1
    out = torch.zeros(N, dtype=torch.float16, device="cuda")
    prog = ops.GemvProgram([dict(weights=[(pd, sd)], x=dev(x), outs=[out])], "q4f16_0")
    prog.run(); torch.cuda.synchronize()
    y = host(out).astype(np.float64)
    base = np.round(np.median(y))
    hits = [(n, y[n] - base) for n in range(N) if abs(y[n] - base) > 0.5]
    print(f"k0={k0:3d}: base={base:7.1f} (n_hits_est={(base*16+1136)/1136:.2f}) special cols: {hits[:8]}")
