"""Where do the fused one-launch program and the reference-segmented programs first differ (bitwise)?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200.decode_stack import DecodeStack

dev = torch.device("cuda", 0)
L = int(sys.argv[1]) if len(sys.argv) > 1 else 4
a = DecodeStack("llama-7b", "q4f16_0", device=dev, layers=L, use_program=True)
b = DecodeStack("llama-7b", "q4f16_0", device=dev, layers=L, use_program=True, segments="reference", weights_from=a)
c = DecodeStack("llama-7b", "q4f16_0", device=dev, layers=L, use_program=True, weights_from=a)
x = (torch.randn(4096, device=dev, generator=torch.Generator(device=dev).manual_seed(5)) * 0.05).half()
for s in (a, b, c):
    s.x_in.copy_(x)
    s.capture()
    s.decode_step()
torch.cuda.synchronize()
for name, other in (("segmented", b), ("second fused stack", c)):
    first = None
    for li in range(L):
        for k in ("q", "k", "v", "x1", "g", "u", "x2"):
            d = (a.buf[li][k].float() - other.buf[li][k].float()).abs()
            n = int((a.buf[li][k] != other.buf[li][k]).sum().item())
            if n and first is None:
                first = (li, k, n, d.max().item(), a.buf[li][k].float().abs().max().item())
    dl = (a.logits - other.logits).abs().max().item()
    print("DIAG finite:", bool(torch.isfinite(a.logits).all()), "max|x2 last|", a.buf[L - 1]["x2"].float().abs().max().item())
    print(f"DIAG {name}: first differing buffer {first}; logits max|diff| {dl:.3e} (max|logit| {a.logits.abs().max().item():.3e})", flush=True)
for _ in range(3):
    y0 = a.decode_step().clone()
    y1 = a.decode_step().clone()
    print("DIAG replay equal:", torch.equal(y0, y1), flush=True)
