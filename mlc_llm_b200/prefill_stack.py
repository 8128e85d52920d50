"""Layer-sequence driver for PREFILL: the quantized-Linear GEMMs of one `encoding` call, in the reference's
order with the reference's fused epilogues, standing in for the compiled Relax `encoding` function
(mlc_llm/relax_model/llama.py:548-578, gpt_neox.py:134-288; call stack SURVEY.md 3.3).

On CUDA the reference runs every Linear of a prefill as TWO kernels - `decode*` re-materialises the whole f16
weight transposed in HBM, then a SIMT `NT_matmul*` with the epilogue fused (quantization.py:372-394,
dispatch/llama/main.py:3984-4232 + 2832-3981, dispatch/gpt_neox/dolly_v2_3b_mod.py:26-275).  Here every Linear is
ONE tcgen05 GEMM that dequantizes inside (`ops.gemm` -> b200q_gemm).

Per layer (M = number of prompt tokens):
  Llama    q|k|v  [M,h]->[M,h] x3 ; o  +residual ; gate  silu ; up  *gate ; down  +residual
           (NT_matmul1, fused_NT_matmul1_add3 / fused_NT_matmul4_add3 (residual), fused_NT_matmul3_silu1,
            fused_NT_matmul3_multiply1; main.py:2832-3660)
  GPT-NeoX q|k|v  +bias (fused_NT_matmul1_add3) ; h_to_4h  +bias, GELU (fused_NT_matmul3_add4_gelu1) ;
           4h_to_h  +bias (fused_NT_matmul4_add3) ; dense  +bias +mlp_out +residual
           (fused_NT_matmul1_add3_add5_add5: the parallel-residual sum of dolly / RedPajama)
Attention, norms and rotary embedding are not on this path: the attention output is stood in for by the q
projection so that shapes and dependencies are the real ones.

Tensor parallel: q/k/v/gate/up/h_to_4h column-sharded, o/down/dense/4h_to_h row-sharded; the
row-sharded GEMMs produce per-rank partial sums [M, h] that are summed with ONE NCCL all-reduce per sharded
block (rank 0 carries bias / residual into the sum so the reduced tensor is the final value).
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .decode_stack import DecodeStack


class PrefillStack:
    """Runs the prefill GEMM sequence on the weights of a DecodeStack (same packed tensors: decode and prefill
    share the weights exactly as in the reference, where both `decoding` and `encoding` read the same params)."""

    def __init__(self, stack: DecodeStack, M: int, fuse_qkv: bool = True, pdl: bool = True):
        if stack.spec.dtype != "float16":
            raise NotImplementedError("prefill GEMM: f16 schemes (b200q_gemm)")
        self.st, self.M = stack, M
        self.fuse_qkv = fuse_qkv and stack.spec.sym
        self.pdl = pdl  # programmatic dependent launch between the GEMMs of the pass (B200Q_FLAG_PDL)
        sh, tp, dev = stack.shape, stack.tp_size, stack.device
        h, i = sh.hidden, sh.inter
        e = lambda *s: torch.zeros(*s, dtype=torch.float16, device=dev)
        self.x_in = e(M, h)
        # one set of activation buffers is reused by every layer (the reference's memory planner does the same)
        self.q, self.k, self.v = e(M, h // tp), e(M, h // tp), e(M, h // tp)
        self.x1, self.x2, self.x3 = e(M, h), e(M, h), e(M, h)
        self.g, self.a = e(M, i // tp), e(M, i // tp)
        self.m_out = e(M, h)
        self.gemms_per_pass = 0
        self.launches_per_pass = 0
        self.collectives_per_pass = 0

    def flops(self) -> int:
        """Algorithmic FLOPs of one pass on THIS rank: 2 M K N per Linear, layers only."""
        tot = 0
        for ws in self.st.layers:
            for w in ws.values():
                tot += 2 * self.M * w.K * w.N
        return tot

    def _gemm(self, w, X, ep, out, ep0=None, ep1=None, ep2=None):
        ops.gemm(w.packed, w.scale, X, self.st.spec, ep, ep0, ep1, ep2, out=out, pdl=self.pdl)
        self.gemms_per_pass += 1
        self.launches_per_pass += 1

    def _qkv(self, w, x, ep, b):
        """q | k | v read the same x: one launch (b200q_gemm_multi), three outputs with their own bias."""
        if self.fuse_qkv:
            ops.gemm_multi([(w[n].packed, w[n].scale) for n in "qkv"], x, self.st.spec, [ep] * 3,
                           [b[n] if b is not None else None for n in "qkv"], outs=[self.q, self.k, self.v], pdl=self.pdl)
            self.gemms_per_pass += 3  # Linears; one launch
            self.launches_per_pass += 1
        else:
            for n, o in (("q", self.q), ("k", self.k), ("v", self.v)):
                self._gemm(w[n], x, ep, o, b[n] if b is not None else None)

    def _row(self, w, X, out, ep, ep0=None, ep1=None, ep2=None):
        """Row-parallel GEMM: partial sums, then one all-reduce per sharded block."""
        st = self.st
        if st.tp_size == 1 or st.tp_rank == 0:
            self._gemm(w, X, ep, out, ep0, ep1, ep2)
        else:
            self._gemm(w, X, "none", out)
        if st.tp_size > 1:
            import torch.distributed as dist

            dist.all_reduce(out, group=st.pg)
            self.collectives_per_pass += 1

    def run(self, x: Optional[torch.Tensor] = None) -> torch.Tensor:
        st = self.st
        self.gemms_per_pass = 0
        self.launches_per_pass = 0
        self.collectives_per_pass = 0
        x = self.x_in if x is None else x
        llama = st.shape.family == "llama"
        for li in range(st.n_layers):
            w = st.layers[li]
            out = self.x2 if (li % 2 == 0) else self.x3  # ping-pong the layer output (x_in is never written)
            if llama:
                self._qkv(w, x, "none", None)
                self._row(w["o"], self.q, self.x1, "add", x)
                self._gemm(w["gate"], self.x1, "silu", self.g)
                self._gemm(w["up"], self.x1, "multiply", self.a, self.g)
                self._row(w["down"], self.a, out, "add", self.x1)
            else:
                b = st.biases[li]
                if self.fuse_qkv and self.M > 128:
                    # q | k | v | h_to_4h all read the layer input (parallel residual): ONE launch, four outputs
                    names = ("q", "k", "v", "h_to_4h")
                    ops.gemm_multi([(w[n].packed, w[n].scale) for n in names], x, self.st.spec, ["bias"] * 3 + ["bias_gelu"],
                                   [b[n] for n in names], outs=[self.q, self.k, self.v, self.a], pdl=self.pdl)
                    self.gemms_per_pass += 4
                    self.launches_per_pass += 1
                else:
                    self._qkv(w, x, "bias", b)
                    self._gemm(w["h_to_4h"], x, "bias_gelu", self.a, b["h_to_4h"])
                self._row(w["4h_to_h"], self.a, self.m_out, "bias", b["4h_to_h"])
                self._row(w["dense"], self.q, out, "bias_add2", b["dense"], self.m_out, x)
            x = out
        return x
