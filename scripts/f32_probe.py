"""q4f32_0 (config 0, RedPajama-3B shapes): tokens/s of the decode program against the per-operator PDL chain, and single GEMVs"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mlc_llm_b200 import ops
from mlc_llm_b200.decode_stack import DecodeStack, token_bytes
from mlc_llm_b200.quant_spec import SPECS
dev = torch.device("cuda", 0)
spec = SPECS["q4f32_0"]
for prog in (True, False):
    st = DecodeStack("redpajama-3b", spec, device=dev, use_program=prog)
    st.x_in.copy_((torch.randn(st.shape.hidden, device=dev) * 0.05).to(st.dt))
    st.capture()
    for _ in range(3): st.decode_step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): st.decode_step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    nb = token_bytes(st.shape, spec, 1)
    print(f"F32 program={prog}: {1000 / ms:.1f} tok/s {ms * 1e3:.0f} us {nb / ms / 1e6:.0f} GB/s launches {st.launches_per_step}", flush=True)
    del st; torch.cuda.empty_cache()
for K, N in ((2560, 2560), (2560, 10240), (10240, 2560), (2560, 50432)):
    p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
    sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).float()
    x = (torch.randn(K, device=dev) * 0.05).float()
    out = torch.empty(N, dtype=torch.float32, device=dev)
    for _ in range(3): ops.gemv(p, sc, x, "q4f32_0", out=out)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50): ops.gemv(p, sc, x, "q4f32_0", out=out)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 50
    nb = spec.gemv_bytes(K, N)
    print(f"F32 gemv {K}x{N}: {us:.1f} us {nb / us / 1e3:.0f} GB/s", flush=True)
