"""The caller side of the path: which tensors are quantized and which operator each call site gets.

Mirrors the reference's GroupQuantize graph rewrite (mlc_llm/transform/quantization.py:362-420):
  * a 2-D parameter W[N, K] used as matmul(x, permute_dims(W)) - every nn.Linear of the model
    (relax_model/llama.py:61-70) - is stored "transposed" as packed uint32[n_words(K), N] + scale[n_groups(K), N]
    (+ min for the f16 asymmetric scheme) and decoded on the fly (:362-395);
  * a call with x.shape[-2] == 1 (the decode step) gets the fused dequant-GEMV, any other call the prefill route
    (:372 `transpose_output = x.struct_info.shape[-2] != 1`; FuseDecodeMatmulEwise, decode_matmul_ewise.py:8-46);
  * a 2-D parameter used by take(W, ids, axis=0) - the token embedding - is stored row-major
    (V, n_words) and gathered with take_decode (:397-420).
`QuantLinear` / `QuantEmbedding` are the torch modules standing in for those call sites; `quantize_model` is the
rewrite (every torch.nn.Linear / torch.nn.Embedding of a module tree), `from_param_cache` builds them from a
reference artefact directory (param_cache.load + group_quantized).  All arithmetic is libb200q's.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .quant_spec import SPECS, QuantSpec


class QuantLinear(torch.nn.Module):
    """y = x @ decode(packed, scale[, min]) (+ bias), epilogue fused.  x: dtype[..., M, K]."""

    def __init__(self, packed, scale, spec, K: int, zmin=None, bias: Optional[torch.Tensor] = None):
        super().__init__()
        self.spec: QuantSpec = SPECS[spec] if isinstance(spec, str) else spec
        self.K, self.N = K, packed.shape[1]
        self.register_buffer("packed", packed)
        self.register_buffer("scale", scale)
        self.register_buffer("zmin", zmin)
        self.register_buffer("bias", bias)

    @classmethod
    def from_linear(cls, lin: torch.nn.Linear, spec):
        """GroupQuantize on one nn.Linear: encode W[N, K] with the library encoder (bit-exact te_encode_*)."""
        spec = SPECS[spec] if isinstance(spec, str) else spec
        W = lin.weight.detach().to(spec.torch_dtype).contiguous()
        r = ops.encode(W, spec, transpose=True)
        bias = None if lin.bias is None else lin.bias.detach().to(spec.torch_dtype).contiguous()
        return cls(r[0], r[1], spec, W.shape[1], zmin=r[2] if spec.has_min else None, bias=bias)

    def forward(self, x: torch.Tensor, epilogue: Optional[str] = None, operand: Optional[torch.Tensor] = None) -> torch.Tensor:
        """epilogue: None -> "bias" if the layer has one else "none"; or any fused epilogue of the path with its
        operand (residual for "add", the gate for "multiply")."""
        ep = epilogue or ("bias" if self.bias is not None else "none")
        e0 = self.bias if ep in ("bias", "bias_gelu") else operand
        x = x.contiguous()
        lead = x.shape[:-1]
        M = x.numel() // self.K
        if M == 1:  # the reference's rule: fuse decode into the matmul only for the constant 1
            odt = torch.float32 if (ep == "cast_f32" or self.spec.dtype == "float32") else torch.float16
            out = torch.empty(lead + (self.N,), dtype=odt, device=x.device)
            return ops.gemv(self.packed, self.scale, x.reshape(1, 1, self.K), self.spec, ep, e0, out=out.view(1, 1, self.N),
                            zmin=self.zmin).view(lead + (self.N,))
        if self.spec.dtype != "float16":
            # f32 prefill is a miss of the dispatch table (as in the reference: decode + a generic f32 matmul)
            Wd = ops.dequantize(self.packed, self.scale, self.K, self.spec, False, zmin=self.zmin)
            y = x.reshape(M, self.K) @ Wd
            y = y + self.bias if self.bias is not None and ep == "bias" else y
            return y.view(lead + (self.N,))
        return ops.gemm(self.packed, self.scale, x.reshape(M, self.K), self.spec, ep, e0, zmin=self.zmin).view(lead + (self.N,))


class QuantEmbedding(torch.nn.Module):
    """take_decode: rows of the quantized table (V, n_words) gathered and decoded."""

    def __init__(self, packed_rm, scale_rm, spec, K: int, zmin_rm=None):
        super().__init__()
        self.spec: QuantSpec = SPECS[spec] if isinstance(spec, str) else spec
        self.K = K
        self.register_buffer("packed_rm", packed_rm)
        self.register_buffer("scale_rm", scale_rm)
        self.register_buffer("zmin_rm", zmin_rm)

    @classmethod
    def from_embedding(cls, emb: torch.nn.Embedding, spec):
        spec = SPECS[spec] if isinstance(spec, str) else spec
        W = emb.weight.detach().to(spec.torch_dtype).contiguous()
        r = ops.encode(W, spec, transpose=False)
        return cls(r[0], r[1], spec, W.shape[1], zmin_rm=r[2] if spec.has_min else None)

    def forward(self, ids: torch.Tensor) -> torch.Tensor:
        flat = ids.reshape(-1).to(torch.int32).contiguous()
        out = ops.take_decode(self.packed_rm, self.scale_rm, flat, self.K, self.spec, zmin_rm=self.zmin_rm)
        return out.view(tuple(ids.shape) + (self.K,))


def quantize_model(module: torch.nn.Module, spec, skip=()) -> torch.nn.Module:
    """The GroupQuantize rewrite on a torch module tree: every nn.Linear -> QuantLinear, every nn.Embedding ->
    QuantEmbedding (in place; `skip`: attribute names to leave alone).  K must suit the scheme (asymmetric: a
    multiple of the group size, quantization.py:107)."""
    for name, child in list(module.named_children()):
        if name in skip:
            continue
        if isinstance(child, torch.nn.Linear):
            setattr(module, name, QuantLinear.from_linear(child, spec))
        elif isinstance(child, torch.nn.Embedding):
            setattr(module, name, QuantEmbedding.from_embedding(child, spec))
        else:
            quantize_model(child, spec, skip)
    return module


def from_param_cache(cache_dir: str, spec, device="cuda"):
    """A reference artefact directory (params/ndarray-cache.json + shards) -> the modules of its quantized
    parameters, in parameter order: [QuantLinear | QuantEmbedding | plain tensor, ...]."""
    from . import param_cache

    spec = SPECS[spec] if isinstance(spec, str) else spec
    params, _ = param_cache.load(cache_dir, device=device)
    mods = []
    for rec in param_cache.group_quantized(params, spec):
        if rec[0] == "linear":
            _, p, s, m, K, N = rec
            mods.append(QuantLinear(p, s, spec, K, zmin=m))
        elif rec[0] == "embedding":
            _, p, s, m, V, K = rec
            mods.append(QuantEmbedding(p, s, spec, K, zmin_rm=m))
        else:
            mods.append(rec[1])
    return mods
