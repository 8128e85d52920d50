"""Loader for the reference's own GPU kernels built by oracle/build_ref_cuda.py  (TEST / BENCH INFRASTRUCTURE ONLY).

oracle/_ref/cuda/<name>.cubin holds the CUDA that TVM generates from the reference's hand-scheduled PrimFuncs
(mlc_llm/dispatch/...), compiled for sm_100a; manifest.json holds each kernel's launch geometry and argument
order.  The kernels are launched through the CUDA driver API (cuda-python) on torch's current stream with torch
tensors as arguments - in the PrimFunc's own parameter order, like R.call_tir does.  Used as "the bar to beat on
the same box" (bench.py `reference_gpu`) and, in tests, as one more check of the oracle against reference code.
"""
from __future__ import annotations

import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
DIR = os.path.join(HERE, "_ref", "cuda")


def available() -> bool:
    return os.path.exists(os.path.join(DIR, "manifest.json"))


def manifest():
    return {m["name"]: m for m in json.load(open(os.path.join(DIR, "manifest.json")))}


class RefKernel:
    def __init__(self, name: str):
        from cuda.bindings import driver as drv

        self.drv = drv
        self.m = manifest()[name]
        data = open(os.path.join(DIR, self.m["cubin"]), "rb").read()
        err, self.mod = drv.cuModuleLoadData(data)
        if int(err) != 0:
            raise RuntimeError(f"cuModuleLoadData({name}): {err}")
        err, self.fn = drv.cuModuleGetFunction(self.mod, self.m["kernel"].encode())
        if int(err) != 0:
            raise RuntimeError(f"cuModuleGetFunction({self.m['kernel']}): {err}")

    def __call__(self, *tensors):
        """tensors in the PrimFunc's parameter order (torch CUDA tensors); launches on torch's current stream."""
        import torch

        vals, types = [], []
        for a in self.m["arg_order"]:
            if isinstance(a, int):
                vals.append(tensors[a].data_ptr())
                types.append(np.uint64)
            else:
                vals.append(int(a.split(":")[1]))
                types.append(np.int32)
        argv = [np.array([v], dtype=t) for v, t in zip(vals, types)]
        ptrs = np.array([a.ctypes.data for a in argv], dtype=np.uint64)
        g, b = self.m["grid"], self.m["block"]
        (err,) = self.drv.cuLaunchKernel(self.fn, g[0], g[1], g[2], b[0], b[1], b[2], 0,
                                         torch.cuda.current_stream().cuda_stream, ptrs.ctypes.data, 0)
        if int(err) != 0:
            raise RuntimeError(f"cuLaunchKernel({self.m['name']}): {err}")
        self._keep = argv  # kernel parameters are copied at launch; keep anyway until the next call
