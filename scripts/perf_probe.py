"""Quick probe of the decode-program kernel (whatever build B200Q_LIB selects): tokens/s of the Llama-7B q4f16_0
token and the bubble-free streaming rate.  python scripts/perf_probe.py [--iters N] [--no-token] [--no-stream]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from mlc_llm_b200.decode_stack import DecodeStack, token_bytes
from mlc_llm_b200.quant_spec import SPECS

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=50)
ap.add_argument("--no-token", action="store_true")
ap.add_argument("--no-stream", action="store_true")
ap.add_argument("--big", action="store_true", help="one 4096 -> 262144 GEMV: steady-state streaming, whole panels per CTA")
ap.add_argument("--model", default="llama-7b")
ap.add_argument("--segments", default="fused")
ap.add_argument("--fmt", default="q4f16_0")
ap.add_argument("--big-n", type=int, default=262144, help="columns of the --big GEMV (16384: 37 MB of q4 weights, L2-resident: the consumer's ceiling)")
a = ap.parse_args()
dev = torch.device("cuda", 0)
spec = SPECS[a.fmt]
tag = os.path.basename(os.environ.get("B200Q_LIB", "libb200q.so"))
if not a.no_token:
    st = DecodeStack(a.model, spec, device=dev, use_program=True, segments=a.segments)
    st.x_in.copy_((torch.randn(st.shape.hidden, device=dev) * 0.05).half())
    st.capture()
    for _ in range(3):
        st.decode_step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        st.decode_step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    nb = token_bytes(st.shape, spec, 1)
    print(f"PROBE {tag} {a.fmt} token {a.model} segments={a.segments}: {1000 / ms:.1f} tok/s  {ms * 1e3:.1f} us  {nb / ms / 1e6:.0f} GB/s "
          f"({nb / ms / 1e6 / 6550.4 * 100:.1f}% of 6550)", flush=True)
    del st
    torch.cuda.empty_cache()
if not a.no_stream:
    r = bench.stream_rate(spec, torch, dev, iters=max(3, a.iters // 5))
    print(f"PROBE {tag} stream: {r['gbs']} GB/s ({r['frac_of_peak'] * 100:.1f}%)", flush=True)

if a.big:
    K, N = 4096, a.big_n
    from mlc_llm_b200 import ops
    x = (torch.randn(K, device=dev) * 0.05).half()
    p = torch.randint(-2**31, 2**31 - 1, (spec.n_words(K), N), dtype=torch.int32, device=dev)
    sc = (torch.rand((spec.n_groups(K), N), device=dev) * 0.01).half()
    out = torch.empty(N, dtype=torch.float16, device=dev)
    prog = ops.GemvProgram([dict(weights=[(p, sc)], x=x, outs=[out])], a.fmt)
    for _ in range(2):
        prog.run()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        prog.run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 10
    nb = spec.gemv_bytes(K, N)
    print(f"PROBE {tag} big 4096->{N}: {us:.1f} us  {nb / us / 1e3:.0f} GB/s ({nb / us / 1e3 / 6550.4 * 100:.1f}% of 6550)", flush=True)
