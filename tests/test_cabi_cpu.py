"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/b200q.h declares, the ctypes prototypes cover them, and argument validation (which runs
before any CUDA call) returns the documented error codes.  No compute happens here."""
import ctypes
import os
import re

import pytest

from mlc_llm_b200 import _lib
from mlc_llm_b200.quant_spec import SPECS, spec_from_flags
from oracle import quant_oracle as qo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "b200q.h")).read()
    return sorted(set(re.findall(r"B200Q_API[^;(]*?\b(b200q_\w+)\s*\(", txt)))


def test_header_symbols_exported_and_bound():
    names = declared_symbols()
    assert len(names) >= 15
    L = _lib.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200q.h but not exported by libb200q.so"
        assert n in _lib.PROTOTYPES, f"{n} has no ctypes prototype"
    assert set(_lib.PROTOTYPES) == set(names)


def test_version_and_layout_helpers_match_oracle():
    L = _lib.lib()
    assert L.b200q_version() == 109
    for name, spec in SPECS.items():
        f = qo.ALL_FORMATS[name]
        assert spec.fmt_id == {"q4f16_0": 0, "q3f16_0": 1, "q4f32_0": 2, "q8f16_0": 3, "q4f16_asym": 4, "q4f32_asym": 5,
                               "q3f16_u16": 6, "fp4f16_0": 7, "fp4f32_0": 8}[name]
        for K in (32, 40, 100, 2560, 4096, 5120, 11008, 13824, 22016):
            assert L.b200q_n_groups(spec.fmt_id, K) == f.n_groups(K) == spec.n_groups(K)
            assert L.b200q_n_words(spec.fmt_id, K) == f.n_words(K) == spec.n_words(K)
        for ep_name, ep in qo.EPILOGUES.items():
            if ep_name == "bias_add2":
                continue
            for K, N in ((4096, 4096), (4096, 11008), (11008, 4096), (4096, 32000)):
                assert L.b200q_gemv_bytes(spec.fmt_id, K, N, ep) == qo.gemv_bytes(K, N, f, ep) == spec.gemv_bytes(K, N, ep_name)
    assert L.b200q_n_groups(9, 64) < 0


def test_epilogue_ids_match():
    from mlc_llm_b200.quant_spec import EPILOGUES

    assert EPILOGUES == qo.EPILOGUES


def test_argument_validation_codes():
    """Validation happens before any CUDA call, so fake (aligned) addresses are enough."""
    L = _lib.lib()
    a = 0x10000  # 16-byte aligned fake device address
    st = None
    # N not a multiple of 8
    assert L.b200q_gemv(a, a, a, None, a, 4096, 12, 0, 0, 0, None, 0, st) == -1
    assert b"multiple of 8" in L.b200q_last_error()
    # unknown format / epilogue
    assert L.b200q_gemv(a, a, a, None, a, 4096, 4096, 7, 0, 0, None, 0, st) == -3
    assert L.b200q_gemv(a, a, a, None, a, 4096, 4096, 0, 42, 0, None, 0, st) == -3
    # epilogue that needs an operand
    assert L.b200q_gemv(a, a, a, None, a, 4096, 4096, 0, 1, 0, None, 0, st) == -5
    # NULL / misaligned
    assert L.b200q_gemv(None, a, a, None, a, 4096, 4096, 0, 0, 0, None, 0, st) == -5
    assert L.b200q_gemv(a + 4, a, a, None, a, 4096, 4096, 0, 0, 0, None, 0, st) == -2
    assert L.b200q_gemv(a, a, a, None, a, 4096, 4096, 0, 0, 0, a + 8, 1 << 20, st) == -2  # workspace
    assert L.b200q_gemv_workspace_bytes() >= 1 << 20
    # tir.noalias: an output that overlaps the activation / an epilogue operand / another output is rejected
    b, c = 0x200000, 0x300000
    assert L.b200q_gemv(a, a, b, None, b, 4096, 4096, 0, 0, 0, None, 0, st) == -3
    assert b"aliases" in L.b200q_last_error()
    assert L.b200q_gemv(a, a, b, c, c, 4096, 4096, 0, 1, 0, None, 0, st) == -3
    assert L.b200q_gemv(a, a, b, None, b + 8192 - 16, 4096, 4096, 0, 0, 0, None, 0, st) == -3  # partial overlap
    assert L.b200q_dequant(a, a, a, 4096, 4100, 0, 0, st) == -1
    assert L.b200q_gemm(a, a, a, None, None, None, a, 16, 4096, 4096, 2, 0, 0, st) == -3  # f32 GEMM
    assert L.b200q_take_decode(a, a, a, a, 0, 10, 64, 0, st) == 0  # empty gather is a no-op
    # asymmetric schemes go through the *_asym entry points (and only they)
    assert L.b200q_gemv(a, a, b, None, c, 4096, 4096, 4, 0, 0, None, 0, st) == -3
    assert L.b200q_gemv_asym(a, a, a, b, None, c, 4096, 4096, 0, 0, 0, None, 0, st) == -3
    assert L.b200q_gemv_asym(a, a, None, b, None, c, 4096, 4096, 4, 0, 0, None, 0, st) == -5  # f16 scheme needs min
    assert L.b200q_gemv_asym(a, a, a, b, None, c, 4096, 4096, 5, 0, 0, None, 0, st) == -3     # f32 scheme packs it
    assert L.b200q_gemv_asym(a, a, a, b, None, c, 4100, 4096, 4, 0, 0, None, 0, st) == -1     # K % 32
    assert L.b200q_gemm_asym(a, a, None, a, None, None, None, c, 16, 4096, 4096, 5, 0, 0, st) == -3  # f32 prefill


def test_round2_entry_points_validate_before_launching():
    """b200q_gemm_ws / b200q_gemm_multi / b200q_nt_matmul_ws / b200q_gemv_static: every refusal below happens before
    any CUDA call, so it can be checked without a GPU."""
    import ctypes

    L = _lib.lib()
    a, b, c, d = 0x100000, 0x4000000, 0x8000000, 0xC000000
    st = None
    # stream-K workspace: alignment; formats without a prefill GEMM
    assert L.b200q_gemm_ws(a, a, None, a, None, None, None, b, 256, 4096, 4096, 0, 0, 0, c + 8, 1 << 20, st) == -2
    assert L.b200q_gemm_ws(a, a, None, a, None, None, None, b, 256, 4096, 4096, 2, 0, 0, None, 0, st) == -3   # q4f32_0
    assert L.b200q_gemm_ws(a, a, None, a, None, None, None, b, 256, 4096, 4096, 6, 0, 0, None, 0, st) == -3   # uint16 int3
    assert L.b200q_gemm_ws(a, a, None, a, None, None, None, b, 256, 4096, 4096, 4, 0, 0, None, 0, st) == -5   # asym needs min
    assert L.b200q_nt_matmul_ws(a, a, None, None, None, b, 256, 4096, 4096, 0, 0, c + 4, 64, st) == -2
    assert L.b200q_nt_matmul_ws(a, a, None, None, None, b, 256, 4100, 4096, 0, 0, None, 0, st) == -1
    # several weights in one launch
    ops4 = (_lib.GemmOp * 5)()
    for i in range(5):
        ops4[i] = _lib.GemmOp(a, a, None, None, None, None, b + i * 0x400000, 512, 0)
    assert L.b200q_gemm_multi(ops4, 5, c, 256, 4096, 0, 0, None, 0, st) == -1 and b"[1,4]" in L.b200q_last_error()
    assert L.b200q_gemm_multi(ops4, 0, c, 256, 4096, 0, 0, None, 0, st) == -1
    assert L.b200q_gemm_multi(None, 2, c, 256, 4096, 0, 0, None, 0, st) == -5
    assert L.b200q_gemm_multi(ops4, 2, c, 0, 4096, 0, 0, None, 0, st) == 0      # M = 0: nothing to do
    ops4[1].N = 12
    assert L.b200q_gemm_multi(ops4, 2, c, 256, 4096, 0, 0, None, 0, st) == -1
    ops4[1].N = 512
    ops4[1].Y = ops4[0].Y + 1024                                                   # outputs overlap
    assert L.b200q_gemm_multi(ops4, 2, c, 256, 4096, 0, 0, None, 0, st) == -3 and b"overlap" in L.b200q_last_error()
    ops4[1].Y = c                                                                  # output aliases X
    assert L.b200q_gemm_multi(ops4, 2, c, 256, 4096, 0, 0, None, 0, st) == -3 and b"aliases X" in L.b200q_last_error()
    ops4[1].Y = b + 0x400000
    ops4[1].epilogue, ops4[1].ep0 = 4, ops4[0].Y                                   # op 1's bias lives in op 0's output
    assert L.b200q_gemm_multi(ops4, 2, c, 256, 4096, 0, 0, None, 0, st) == -3
    assert L.b200q_gemm_multi(ops4, 2, c, 256, 4096, 2, 0, None, 0, st) == -3      # f32 format
    # the allocation-free GEMV: formats without a split-K kernel, bad shapes
    assert L.b200q_gemv_static(a, a, b, None, c, 40, 8, 6, 0, 0, st) == -3
    assert L.b200q_gemv_static(a, a, b, None, c, 4096, 12, 0, 0, 0, st) == -1
    assert L.b200q_gemv_static(a, a, b, None, b, 4096, 4096, 0, 0, 0, st) == -3    # out aliases x
    # host-side schedule query
    out = (ctypes.c_int * 20)()
    assert L.b200q_debug_gemm_schedule(1, 80, 160, 4, 74, 0, out, 4) == 2


def test_spec_from_flags():
    assert spec_from_flags("float16", "int4", True, 32).name == "q4f16_0"
    assert spec_from_flags("float16", "int3", True, 32).name == "q3f16_0"
    assert spec_from_flags("float32", "int4", True, 32).name == "q4f32_0"
    assert spec_from_flags("float16", "int8", True, 32).name == "q8f16_0"
    # the reference CLI's default (build.py:19: --quantization-sym off) is the asymmetric scheme
    assert spec_from_flags("float16", "int4", False, 32).name == "q4f16_asym"
    assert spec_from_flags("float32", "int4", False, 32).name == "q4f32_asym"
    with pytest.raises(NotImplementedError):
        spec_from_flags("float16", "int3", False, 32)  # quantization.py:283-284: int3 asserts sym
    assert spec_from_flags("float16", "int3", True, 16).name == "q3f16_u16"  # the iOS build (ios/README.md:39)
    assert spec_from_flags("float16", "fp4", True, 32).name == "fp4f16_0"
    assert spec_from_flags("float32", "fp4", True, 32).name == "fp4f32_0"
    with pytest.raises(NotImplementedError):
        spec_from_flags("float16", "int4", True, 8)


def test_ops_refuse_cpu_tensors():
    import torch

    from mlc_llm_b200 import ops

    p = torch.zeros((512, 64), dtype=torch.int32)
    s = torch.zeros((128, 64), dtype=torch.float16)
    x = torch.zeros((1, 1, 4096), dtype=torch.float16)
    with pytest.raises(ValueError, match="no CPU path"):
        ops.gemv(p, s, x, "q4f16_0")
