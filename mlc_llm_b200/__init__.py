"""mlc_llm_b200 - B200-native (sm_100a) replacement for ONE hot path of mlc-llm: the fused
dequantize + matmul over the q3f16_0 / q4f16_0 / q4f32_0 / q8f16_0 packed-uint32 weight layouts
(decode-step GEMV, prefill dequantize / GEMM, tensor-parallel sharding of both).

Everything computes in hand-written CUDA behind the C-ABI of include/b200q.h; this package is the
Python host side (the reference's host side is Python too): ctypes binding, torch-tensor wrappers,
the dispatch registry mirroring mlc_llm/dispatch, the operator factories mirroring
mlc_llm/transform/quantization.py, and the layer-sequence / tensor-parallel drivers.
"""
from .quant_spec import EPILOGUES, SPECS, QuantSpec, spec_from_flags  # noqa: F401

__version__ = "0.1.0"
