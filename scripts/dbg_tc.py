"""Debug driver for the tcgen05 program kernel: structured inputs that expose layout mistakes."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from mlc_llm_b200 import ops
from oracle import quant_oracle as qo
from util import dev, host
fmt = qo.FORMATS["q4f16_0"]
np.set_printoptions(linewidth=200, precision=4, suppress=True)

def run(K, N, q, scale, x):
    # q: [K, N] codes 0..15
    R = K // 8
    packed = np.zeros((R, N), dtype=np.uint32)
    for t in range(8):
        packed |= (q[t::8, :].astype(np.uint32) & 15) << (4 * t)
    out = torch.zeros(N, dtype=torch.float16, device="cuda")
    prog = ops.GemvProgram([dict(weights=[(dev(packed), dev(scale))], x=dev(x), outs=[out])], "q4f16_0")
    prog.run(); torch.cuda.synchronize()
    y = host(out).astype(np.float64)
    w = (q.astype(np.float64) - 7) * np.repeat(scale.astype(np.float64), 32, axis=0)
    ref = x.astype(np.float64) @ w
    return y, ref

K, N = int(sys.argv[1]) if len(sys.argv) > 1 else 128, int(sys.argv[2]) if len(sys.argv) > 2 else 128
rng = np.random.default_rng(0)
ones = np.ones((K // 32, N), dtype=np.float16)
# A: all-zero weights (q = 7), random x
y, ref = run(K, N, np.full((K, N), 7), ones, rng.standard_normal(K).astype(np.float16))
print("A q=7: max|y| =", np.abs(y).max())
# B: one-hot x
q = rng.integers(0, 16, size=(K, N))
for k in [0, 1, 4, 5, 8, 9, 31, 32, 33, 40, 64, 100, K - 1]:
    if k >= K: continue
    x = np.zeros(K, dtype=np.float16); x[k] = 1
    y, ref = run(K, N, q, ones, x)
    bad = np.abs(y - ref) > 1e-3
    print(f"B onehot k={k}: mismatches {bad.sum()}/{N}", "" if not bad.any() else f"first bad n={np.argmax(bad)} y={y[bad][:6]} ref={ref[bad][:6]}")
# C: random everything
x = rng.standard_normal(K).astype(np.float16)
sc = (rng.random((K // 32, N)) * 0.01).astype(np.float16)
y, ref = run(K, N, q, sc, x)
print("C random: max err", np.abs(y - ref).max(), "max|ref|", np.abs(ref).max())
