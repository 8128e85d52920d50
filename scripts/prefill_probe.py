"""config 3 prefill pass (dolly-v2-3b q3f16_0, M = 2048): with / without programmatic dependent launch, fused q|k|v or not"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200.decode_stack import DecodeStack
from mlc_llm_b200.prefill_stack import PrefillStack
dev = torch.device("cuda", 0)
st = DecodeStack("dolly-v2-3b", "q3f16_0", device=dev, use_program=False)
for pdl, fuse in ((False, False), (False, True), (True, True)):
    pf = PrefillStack(st, 2048, fuse_qkv=fuse, pdl=pdl)
    pf.x_in.copy_((torch.randn(2048, st.shape.hidden, device=dev) * 0.05).half())
    for _ in range(3): pf.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): pf.run()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"PREFILL pdl={pdl} fuse_qkv={fuse}: {ms:.3f} ms {pf.flops() / ms / 1e9:.1f} TFLOP/s launches {pf.launches_per_pass}", flush=True)
    y = pf.run().float()
    print("  checksum", float(y.abs().mean()), bool(torch.isfinite(y).all()), flush=True)
