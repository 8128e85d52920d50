"""Small invocations of every kernel family for compute-sanitizer (memcheck / synccheck / racecheck):
the tcgen05 decode program (2 tiny layers), the generic program kernel, the stream GEMV, the tcgen05 GEMM,
dequantize / take_decode / encode.  Results are checked for finiteness only (parity lives in tests/)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.decode_stack import DecodeStack, ModelShape

dev = torch.device("cuda", 0)
shape = ModelShape("tiny-llama", "llama", 512, 1408, 2, 1024)
for fmt in ("q4f16_0", "q8f16_0"):
    st = DecodeStack(shape, fmt, device=dev, layers=2, use_program=True, use_graph=False)
    st.x_in.copy_((torch.randn(shape.hidden, device=dev) * 0.5).half())
    for _ in range(2):
        y = st.decode_step()
    torch.cuda.synchronize()
    assert torch.isfinite(y).all()
    print("program", fmt, "ok", flush=True)
st = DecodeStack(shape, "q3f16_0" if False else "q4f16_0", device=dev, layers=1, use_program=False, use_graph=False)
st.x_in.copy_((torch.randn(shape.hidden, device=dev) * 0.5).half())
y = st.decode_step()
torch.cuda.synchronize()
assert torch.isfinite(y).all()
print("stream ok", flush=True)
w = st.layers[0]["gate"]
X = (torch.randn(256, w.K, device=dev) * 0.1).half()
Y = ops.gemm(w.packed, w.scale, X, "q4f16_0", "silu")
Wd = ops.dequantize(w.packed, w.scale, w.K, "q4f16_0", True)
Y2 = ops.nt_matmul(X, Wd)
p, s = ops.encode(Wd, "q4f16_0")
torch.cuda.synchronize()
assert torch.isfinite(Y).all() and torch.isfinite(Y2).all()
print("gemm/dequant/encode ok", flush=True)
