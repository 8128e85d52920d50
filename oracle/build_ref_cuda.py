"""Recipe: the REFERENCE's own hand-scheduled GPU kernels, compiled for sm_100a  (TEST / BENCH INFRASTRUCTURE ONLY).

SURVEY.md 8(c)/(d): the reference's scheduled PrimFuncs lower to CUDA C through TVM's CUDA backend.  This script
(run in the build container, where /root/reference and the TVM bundled with `tilelang` exist) emits that CUDA C
for the hot-path kernels, records the grid / block sizes of every schedule in a manifest, and compiles each
kernel into oracle/_ref/cuda/<name>.cubin with nvcc for sm_100a (loaded through the driver API by oracle/ref_cuda.py).  Nothing of the reference is
copied into the repository: oracle/_ref/ is git-ignored build output (it travels to the GPU box with the
snapshot).  bench.py's `reference_gpu` record and tests/test_reference_gpu.py load the cubins when present.

Kernels (reference file:line):
  gemv_q4f16_0_4096x4096            sch_fused_decode4_matmul3                 dispatch_tir_operator_adreno.py:25-63
  gemv_q4f16_0_4096x11008_silu      sch_fused_decode5_fused_matmul6_silu1     dispatch_tir_operator_adreno.py:230-272
  gemv_q4f16_0_11008x4096_add       sch_fused_decode6_fused_matmul7_add1      dispatch_tir_operator_adreno.py:302-341  (*)
  gemv_q4f16_0_4096x32000_cast      sch_fused_decode3_fused_matmul1_cast2     dispatch_tir_operator_adreno.py:371-411  (*)
  gemv_q4f16_asym_4096x4096         fused_decode4_matmul5_fp16_after          dispatch/llama/main.py:5000-5053
  decode_q4f16_asym_4096x4096       decode_sch_func(decode4_fp16)             dispatch/llama/main.py:4042-4058, 4206-4232
  nt_matmul_f16_4096x4096           NT_matmul1_fp16_sch_func()                dispatch/llama/main.py:3885-3981   (n = 2048)

(*) these two schedules do not lower with the bundled TVM 0.23 (`target.target_has_feature` is gone: API drift);
    they are attempted and reported as FAILED, the other five build.

Usage:  PYTHONDONTWRITEBYTECODE=1 python oracle/build_ref_cuda.py
"""
from __future__ import annotations

import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(HERE, "_ref", "cuda")
REF = "/root/reference"
sys.path.insert(0, ROOT)

NVCC = "/usr/local/cuda/bin/nvcc"
N_DYN = 2048  # value bound to the dynamic sequence length `n` of the prefill kernels


def thread_extents(tvm, pf, dyn):
    """blockIdx / threadIdx extents of a scheduled PrimFunc (dynamic vars substituted)."""
    from tvm import tir

    ext = {}

    def visit(node):
        if isinstance(node, tir.For) and node.thread_binding is not None:
            tag = str(node.thread_binding.thread_tag)
            e = node.extent
            if dyn:
                e = tir.stmt_functor.substitute(e, dyn)
            e = tvm.arith.Analyzer().simplify(e)
            ext[tag] = max(ext.get(tag, 1), int(e))

    tir.stmt_functor.post_order_visit(pf.body, visit)
    return ext


def emit(tvm, pf, name, dyn_names=()):
    from tvm import tir

    pf = pf.with_attr("global_symbol", name)
    tgt = tvm.target.Target({"kind": "cuda", "arch": "sm_100a", "max_threads_per_block": 1024,
                             "max_shared_memory_per_block": 49152, "thread_warp_size": 32}, host="c")
    lib = tvm.compile(pf, target=tgt)
    mods = lib.mod.imports if hasattr(lib, "mod") else lib.imported_modules
    src = None
    for m in mods:
        src = m.inspect_source() if hasattr(m, "inspect_source") else m.get_source()
    assert src and "__global__" in src, name
    # dynamic vars of the PrimFunc (match_buffer shapes)
    dyn = {}
    for p in pf.params:
        b = pf.buffer_map.get(p)
        if b is None:
            continue
        for s in b.shape:
            if isinstance(s, tir.Var):
                dyn[s] = tir.const(N_DYN, s.dtype)
    ext = thread_extents(tvm, pf, dyn)
    grid = [ext.get(f"blockIdx.{a}", 1) for a in "xyz"]
    block = [ext.get(f"threadIdx.{a}", 1) for a in "xyz"]
    # the kernel's own parameter list decides the launcher's argument order
    m = re.search(r'extern "C" __global__ void (?:__launch_bounds__\(\d+\) )?(\w+)\(([^)]*)\)', src)
    kname, params = m.group(1), [p.strip() for p in m.group(2).split(",")]
    pnames = [p.split()[-1].lstrip("*") for p in params]
    # map kernel parameters to PrimFunc buffers (TVM names the kernel params after the buffers / vars)
    bufs = {pf.buffer_map[p].name: (i, pf.buffer_map[p]) for i, p in enumerate(pf.params) if p in pf.buffer_map}
    order = []
    for pn, ptxt in zip(pnames, params):
        if pn in bufs:
            order.append(("buf", bufs[pn][0]))
        else:
            order.append(("int", pn))  # a dynamic extent such as n
    os.makedirs(OUT, exist_ok=True)
    cu = os.path.join(OUT, name + ".cu")
    open(cu, "w").write(src)
    # device-only compile (TVM's CUDA source is written for NVRTC: its preamble needs __CUDA_ARCH__), loaded with the
    # driver API by oracle/ref_cuda.py
    cubin = os.path.join(OUT, name + ".cubin")
    subprocess.check_call([NVCC, "-O3", "-arch=sm_100a", "-cubin", "-w", "-o", cubin, cu])
    shapes = []
    for k, i in order:
        if k == "buf":
            b = pf.buffer_map[pf.params[i]]
            shapes.append({"param": i, "dtype": str(b.dtype),
                           "shape": [N_DYN if isinstance(s, tir.Var) else int(s) for s in b.shape]})
    return {"name": name, "kernel": kname, "grid": grid, "block": block, "args": shapes, "cubin": os.path.basename(cubin),
            "arg_order": [i if k == "buf" else f"int:{N_DYN}" for k, i in order]}


def main():
    from oracle.gen_golden import _load_tvm

    tvm = _load_tvm()
    sys.path.insert(0, REF)
    import mlc_llm  # noqa: F401
    from mlc_llm.dispatch import dispatch_tir_operator_adreno as ad
    from mlc_llm.dispatch.llama import main as lm

    jobs = [
        ("gemv_q4f16_0_4096x4096", lambda: ad.sch_fused_decode4_matmul3(ad.fused_decode4_matmul3)),
        ("gemv_q4f16_0_4096x11008_silu", lambda: ad.sch_fused_decode5_fused_matmul6_silu1(ad.fused_decode5_fused_matmul6_silu1)),
        ("gemv_q4f16_0_11008x4096_add", lambda: ad.sch_fused_decode6_fused_matmul7_add1(ad.fused_decode6_fused_matmul7_add1)),
        ("gemv_q4f16_0_4096x32000_cast", lambda: ad.sch_fused_decode3_fused_matmul1_cast2(ad.fused_decode3_fused_matmul1_cast2)),
        ("gemv_q4f16_asym_4096x4096", lambda: lm.fused_decode4_matmul5_fp16_after),
        ("decode_q4f16_asym_4096x4096", lambda: lm.decode_sch_func(lm.decode4_fp16)),
        ("nt_matmul_f16_4096x4096", lambda: lm.NT_matmul1_fp16_sch_func()),
    ]
    manifest = []
    for name, get in jobs:
        try:
            rec = emit(tvm, get(), name)
            manifest.append(rec)
            print("built", name, "grid", rec["grid"], "block", rec["block"], [a["shape"] for a in rec["args"]])
        except Exception as e:  # noqa: BLE001
            print("FAILED", name, repr(e)[:300])
    json.dump(manifest, open(os.path.join(OUT, "manifest.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
