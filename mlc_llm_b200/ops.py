"""PyTorch-tensor wrappers over the C-ABI (device memory and streams come from torch; every
computation is a kernel of libb200q.so, launched on torch's current stream).

Argument order follows the reference PrimFuncs: (packed, scale, x[, epilogue operands]) -> out
(mlc_llm/dispatch/dispatch_tir_operator_adreno.py:6,128).  Outputs may be passed in (destination
passing, like R.call_tir) or are allocated here.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from . import _lib
from .quant_spec import EPILOGUES, FLAG_PDL, SPECS, QuantSpec


def _spec(spec) -> QuantSpec:
    return SPECS[spec] if isinstance(spec, str) else spec


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _need_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise ValueError("b200q ops take CUDA tensors only (there is no CPU path)")
        if t is not None and not t.is_contiguous():
            raise ValueError("b200q ops take contiguous tensors (compact row-major, as the reference's buffers)")


_WORKSPACES: dict = {}


def workspace(device=None) -> torch.Tensor:
    """The GEMV split-K scratch of (device, current stream): allocated and zero-filled once, kept
    clean by the kernels (include/b200q.h).  One per stream because concurrent launches must not
    share it."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), _stream())
    ws = _WORKSPACES.get(key)
    if ws is None:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("allocate the b200q workspace before graph capture: call ops.workspace() or run the op once on this stream")
        ws = torch.zeros(int(_lib.lib().b200q_gemv_workspace_bytes()), dtype=torch.uint8, device=dev)
        _WORKSPACES[key] = ws
    return ws


def _ws_args(ws):
    """(pointer, bytes) of a workspace tensor; False = run without one."""
    if ws is False:
        return None, 0
    if ws is None:
        ws = workspace()
    return ws.data_ptr(), ws.numel() * ws.element_size()


def _ep_id(epilogue) -> int:
    return EPILOGUES[epilogue] if isinstance(epilogue, str) else int(epilogue)


def _check_weight(spec: QuantSpec, packed: torch.Tensor, scale: torch.Tensor, K: int):
    ok = (torch.int16, torch.uint16) if spec.storage_nbit == 16 else (torch.int32, torch.uint32)
    if packed.dtype not in ok or packed.dim() != 2:
        raise ValueError(f"packed must be a 2-D uint{spec.storage_nbit} (or int{spec.storage_nbit} view) tensor [n_words, N]")
    N = packed.shape[1]
    if packed.shape[0] != spec.n_words(K):
        raise ValueError(f"packed has {packed.shape[0]} word rows, {spec.name} with K={K} needs {spec.n_words(K)}")
    if scale.dtype != spec.scale_torch_dtype or tuple(scale.shape) != (spec.n_groups(K), N):
        raise ValueError(f"scale must be {spec.scale_torch_dtype}[{spec.n_groups(K)}, {N}]")
    return N


def _check_min(spec: QuantSpec, scale: torch.Tensor, zmin: Optional[torch.Tensor]):
    """The group-minimum tensor of the f16 asymmetric scheme (None for every other scheme)."""
    if spec.has_min:
        if zmin is None or zmin.dtype != scale.dtype or zmin.shape != scale.shape:
            raise ValueError(f"{spec.name}: zmin must be a tensor shaped like scale")
    elif zmin is not None:
        raise ValueError(f"{spec.name} has no separate min tensor")


def gemv(
    packed: torch.Tensor,
    scale: torch.Tensor,
    x: torch.Tensor,
    spec,
    epilogue="none",
    ep0: Optional[torch.Tensor] = None,
    out: Optional[torch.Tensor] = None,
    pdl: bool = False,
    ws=None,
    zmin: Optional[torch.Tensor] = None,
) -> torch.Tensor:
    """fused_decode*_matmul*: y[..., n] = sum_k x[..., k] * decode(packed, scale)[k, n] (+ epilogue).

    x: dtype[..., K] with every leading dim == 1 (the reference fuses only for the constant 1,
    transform/decode_matmul_ewise.py:8-11).  Returns dtype[1, 1, N] (float32 for cast_f32)."""
    spec = _spec(spec)
    ep = _ep_id(epilogue)
    _need_cuda(packed, scale, x, ep0, out)
    K = x.shape[-1]
    if x.numel() != K:
        raise ValueError("gemv is the batch-1 decode path: x must have exactly K elements")
    if x.dtype != spec.torch_dtype:
        raise ValueError(f"x must be {spec.dtype}")
    N = _check_weight(spec, packed, scale, K)
    odt = torch.float32 if (ep == EPILOGUES["cast_f32"] or spec.dtype == "float32") else torch.float16
    if out is None:
        out = torch.empty((1, 1, N), dtype=odt, device=x.device)
    elif out.numel() != N or out.dtype != odt:
        raise ValueError(f"out must be {odt}[{N}]")
    if ep0 is not None and (ep0.numel() != N or ep0.dtype != spec.torch_dtype):
        raise ValueError(f"ep0 must be {spec.dtype}[{N}]")
    L = _lib.lib()
    wp, wb = _ws_args(ws)
    _need_cuda(zmin)
    _check_min(spec, scale, zmin)
    if not spec.sym:  # (packed, scale, min, x[, operand], out): dispatch/llama/main.py:4979
        _lib.check(L.b200q_gemv_asym(packed.data_ptr(), scale.data_ptr(), _ptr(zmin), x.data_ptr(), _ptr(ep0), out.data_ptr(),
                                     K, N, spec.fmt_id, ep, FLAG_PDL if pdl else 0, wp, wb, _stream()))
        return out
    _lib.check(
        L.b200q_gemv(
            packed.data_ptr(), scale.data_ptr(), x.data_ptr(), _ptr(ep0), out.data_ptr(),
            K, N, spec.fmt_id, ep, FLAG_PDL if pdl else 0, wp, wb, _stream(),
        )
    )
    return out


def gemv_static(packed, scale, x, spec, epilogue="none", ep0=None, out=None, pdl: bool = False) -> torch.Tensor:
    """b200q_gemv_static: `gemv` with the split-K scratch in the library's static device array - the entry point itself
    allocates nothing (pass `out` to keep the whole call allocation-free).  Calls that overlap on the GPU race on the
    shared array; N <= 524288.  No counterpart in the reference."""
    spec = _spec(spec)
    ep = _ep_id(epilogue)
    _need_cuda(packed, scale, x, ep0, out)
    K = x.shape[-1]
    if x.numel() != K:
        raise ValueError("gemv is the batch-1 decode path: x must have exactly K elements")
    if x.dtype != spec.torch_dtype:
        raise ValueError(f"x must be {spec.dtype}")
    N = _check_weight(spec, packed, scale, K)
    odt = torch.float32 if (ep == EPILOGUES["cast_f32"] or spec.dtype == "float32") else torch.float16
    if out is None:
        out = torch.empty((1, 1, N), dtype=odt, device=x.device)
    elif out.numel() != N or out.dtype != odt:
        raise ValueError(f"out must be {odt}[{N}]")
    if ep0 is not None and (ep0.numel() != N or ep0.dtype != spec.torch_dtype):
        raise ValueError(f"ep0 must be {spec.dtype}[{N}]")
    _lib.check(_lib.lib().b200q_gemv_static(packed.data_ptr(), scale.data_ptr(), x.data_ptr(), _ptr(ep0), out.data_ptr(),
                                            K, N, spec.fmt_id, ep, FLAG_PDL if pdl else 0, _stream()))
    return out


def gemv_multi(
    weights: Sequence[tuple],
    x: torch.Tensor,
    spec,
    epilogues: Optional[Sequence] = None,
    ep0s: Optional[Sequence[Optional[torch.Tensor]]] = None,
    outs: Optional[Sequence[torch.Tensor]] = None,
    pdl: bool = False,
    x_mul: Optional[torch.Tensor] = None,
    ws=None,
):
    """Up to four GEMVs that share x (q|k|v, gate|up) in one launch.  weights: [(packed, scale), ...].
    x_mul: the activation is x * x_mul (the SwiGLU product read by down_proj)."""
    spec = _spec(spec)
    n = len(weights)
    epilogues = list(epilogues) if epilogues is not None else ["none"] * n
    ep0s = list(ep0s) if ep0s is not None else [None] * n
    K = x.shape[-1]
    _need_cuda(x, x_mul)
    if x.numel() != K or x.dtype != spec.torch_dtype:
        raise ValueError("x must be dtype[K]")
    if x_mul is not None and (x_mul.numel() != K or x_mul.dtype != spec.torch_dtype):
        raise ValueError("x_mul must be dtype[K]")
    ops = (_lib.GemvOp * n)()
    results = []
    for i, (packed, scale) in enumerate(weights):
        _need_cuda(packed, scale, ep0s[i])
        N = _check_weight(spec, packed, scale, K)
        ep = _ep_id(epilogues[i])
        odt = torch.float32 if (ep == EPILOGUES["cast_f32"] or spec.dtype == "float32") else torch.float16
        o = outs[i] if outs is not None else torch.empty((1, 1, N), dtype=odt, device=x.device)
        if o.numel() != N or o.dtype != odt:
            raise ValueError(f"out[{i}] must be {odt}[{N}]")
        ops[i] = _lib.GemvOp(packed.data_ptr(), scale.data_ptr(), _ptr(ep0s[i]), o.data_ptr(), N, ep)
        results.append(o)
    L = _lib.lib()
    wp, wb = _ws_args(ws)
    _lib.check(L.b200q_gemv_multi(ops, n, x.data_ptr(), _ptr(x_mul), K, spec.fmt_id, FLAG_PDL if pdl else 0, wp, wb,
                                  _stream()))
    return results


def gemv_tune(ctas_per_sm: int = 0, stages: int = 0, dbg_flags: int = 0) -> None:
    _lib.lib().b200q_gemv_tune(ctas_per_sm, stages, dbg_flags)


def dequantize(packed, scale, K: int, spec, transpose_output: bool = False, out=None, zmin=None) -> torch.Tensor:
    """decode*: packed/scale -> dtype[K, N] (or [N, K]); bit-exact against the reference decode."""
    spec = _spec(spec)
    _need_cuda(packed, scale, out)
    N = _check_weight(spec, packed, scale, K)
    shape = (N, K) if transpose_output else (K, N)
    if out is None:
        out = torch.empty(shape, dtype=spec.torch_dtype, device=packed.device)
    elif tuple(out.shape) != shape or out.dtype != spec.torch_dtype:
        raise ValueError(f"out must be {spec.dtype}{list(shape)}")
    L = _lib.lib()
    _need_cuda(zmin)
    _check_min(spec, scale, zmin)
    if not spec.sym:
        _lib.check(L.b200q_dequant_asym(packed.data_ptr(), scale.data_ptr(), _ptr(zmin), out.data_ptr(), K, N, spec.fmt_id,
                                        int(transpose_output), _stream()))
        return out
    _lib.check(L.b200q_dequant(packed.data_ptr(), scale.data_ptr(), out.data_ptr(), K, N, spec.fmt_id,
                               int(transpose_output), _stream()))
    return out


def take_decode(packed_rm, scale_rm, ids, K: int, spec, out=None, zmin_rm=None) -> torch.Tensor:
    """take_decode: quantized embedding gather from the row-major (V, n_words) layout."""
    spec = _spec(spec)
    _need_cuda(packed_rm, scale_rm, ids, out)
    V = packed_rm.shape[0]
    if packed_rm.shape[1] != spec.n_words(K) or tuple(scale_rm.shape) != (V, spec.n_groups(K)):
        raise ValueError("packed_rm / scale_rm do not match (V, n_words) / (V, n_groups)")
    if ids.dtype != torch.int32 or ids.dim() != 1:
        raise ValueError("ids must be a 1-D int32 tensor")
    n = ids.numel()
    if out is None:
        out = torch.empty((n, K), dtype=spec.torch_dtype, device=packed_rm.device)
    L = _lib.lib()
    _need_cuda(zmin_rm)
    _check_min(spec, scale_rm, zmin_rm)
    if not spec.sym:
        _lib.check(L.b200q_take_decode_asym(packed_rm.data_ptr(), scale_rm.data_ptr(), _ptr(zmin_rm), ids.data_ptr(),
                                            out.data_ptr(), n, V, K, spec.fmt_id, _stream()))
        return out
    _lib.check(L.b200q_take_decode(packed_rm.data_ptr(), scale_rm.data_ptr(), ids.data_ptr(), out.data_ptr(),
                                   n, V, K, spec.fmt_id, _stream()))
    return out


def encode(W: torch.Tensor, spec, transpose: bool = True):
    """encoding_func (te_encode_sym): W dtype[N, K] -> (packed uint32, scale dtype); bit-exact."""
    spec = _spec(spec)
    _need_cuda(W)
    if W.dim() != 2 or W.dtype != spec.torch_dtype:
        raise ValueError(f"W must be {spec.dtype}[N, K]")
    N, K = W.shape
    R, G = spec.n_words(K), spec.n_groups(K)
    pshape, sshape = ((R, N), (G, N)) if transpose else ((N, R), (N, G))
    packed = torch.empty(pshape, dtype=spec.packed_torch_dtype, device=W.device)
    scale = torch.empty(sshape, dtype=spec.scale_torch_dtype, device=W.device)
    L = _lib.lib()
    if not spec.sym:  # te_encode_asym: (packed, scale, min) for f16, (packed, scale_min) for f32
        if K % spec.group_size:
            raise ValueError("asymmetric encode: K must be a multiple of the group size (quantization.py:107)")
        zmin = torch.empty(sshape, dtype=spec.torch_dtype, device=W.device) if spec.has_min else None
        _lib.check(L.b200q_encode_asym(W.data_ptr(), packed.data_ptr(), scale.data_ptr(), _ptr(zmin), N, K, spec.fmt_id,
                                       int(transpose), _stream()))
        return (packed, scale, zmin) if spec.has_min else (packed, scale)
    _lib.check(L.b200q_encode(W.data_ptr(), packed.data_ptr(), scale.data_ptr(), N, K, spec.fmt_id,
                              int(transpose), _stream()))
    return packed, scale


GEMM_TILES = {None: 0, "single": 2, "pair128": 4, "pair256": 8}  # B200Q_FLAG_GEMM_*


def program_uses_tensor_cores(spec) -> bool:
    """True when b200q_gemv_program_* runs this format on the tcgen05 kernel (group scale applied to the f32 group
    sum: include/b200q.h "arithmetic"): q4f16_0 and q8f16_0, unless B200Q_PROG_TC / B200Q_PROG_TC8 = 0."""
    import os
    name = _spec(spec).name
    if os.environ.get("B200Q_PROG_TC", "1") == "0":
        return False
    return name == "q4f16_0" or (name == "q8f16_0" and os.environ.get("B200Q_PROG_TC8", "1") != "0")


_GEMM_WORKSPACES: dict = {}
FLAG_GEMM_NO_STREAMK = 16
FLAG_GEMM_STREAMK = 32


def gemm_workspace(device=None) -> torch.Tensor:
    """The prefill GEMM's stream-K scratch of (device, current stream): allocated and zero-filled once, left clean by
    every launch (include/b200q.h, b200q_gemm_ws).  One per stream: overlapping launches must not share it."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), _stream())
    ws = _GEMM_WORKSPACES.get(key)
    if ws is None:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("allocate the b200q GEMM workspace before graph capture: call ops.gemm_workspace() first")
        with torch.cuda.device(dev):
            ws = torch.zeros(int(_lib.lib().b200q_gemm_workspace_bytes()), dtype=torch.uint8, device=dev)
        _GEMM_WORKSPACES[key] = ws
    return ws


def gemm(packed, scale, X, spec, epilogue="none", ep0=None, ep1=None, ep2=None, out=None, pdl: bool = False, tile=None,
         zmin=None, ws=None, streamk: bool = True):
    """fused_decode*_NT_matmul* / decode + NT_matmul: Y[m, n] = sum_k X[m, k] * w[k, n] (+ epilogue)
    on the tcgen05 tensor cores.  X dtype[..., M, K] -> Y dtype[..., M, N].
    ws: stream-K workspace (None: the stream's gemm_workspace(); False: b200q_gemm without one); streamk=False keeps
    the whole-tile schedule even with a workspace, streamk="force" cuts the k range whatever the shape."""
    spec = _spec(spec)
    ep = _ep_id(epilogue)
    _need_cuda(packed, scale, X, ep0, ep1, ep2, out)
    K = X.shape[-1]
    M = X.numel() // K
    N = _check_weight(spec, packed, scale, K)
    if X.dtype != spec.torch_dtype:
        raise ValueError(f"X must be {spec.dtype}")
    odt = torch.float32 if ep == EPILOGUES["cast_f32"] else spec.torch_dtype
    oshape = tuple(X.shape[:-1]) + (N,)
    if out is None:
        out = torch.empty(oshape, dtype=odt, device=X.device)
    elif out.numel() != M * N or out.dtype != odt:
        raise ValueError(f"out must be {odt}[{M}, {N}]")
    L = _lib.lib()
    _need_cuda(zmin)
    _check_min(spec, scale, zmin)
    flags = (FLAG_PDL if pdl else 0) | GEMM_TILES[tile]
    flags |= FLAG_GEMM_STREAMK if streamk == "force" else (0 if streamk else FLAG_GEMM_NO_STREAMK)
    if ws is not False:
        w = gemm_workspace(X.device) if ws is None else ws
        _lib.check(L.b200q_gemm_ws(packed.data_ptr(), scale.data_ptr(), _ptr(zmin), X.data_ptr(), _ptr(ep0), _ptr(ep1), _ptr(ep2),
                                   out.data_ptr(), M, K, N, spec.fmt_id, ep, flags, w.data_ptr(), w.numel() * w.element_size(), _stream()))
        return out
    if not spec.sym:
        _lib.check(L.b200q_gemm_asym(packed.data_ptr(), scale.data_ptr(), _ptr(zmin), X.data_ptr(), _ptr(ep0), _ptr(ep1), _ptr(ep2),
                                     out.data_ptr(), M, K, N, spec.fmt_id, ep, flags, _stream()))
        return out
    _lib.check(L.b200q_gemm(packed.data_ptr(), scale.data_ptr(), X.data_ptr(), _ptr(ep0), _ptr(ep1), _ptr(ep2),
                            out.data_ptr(), M, K, N, spec.fmt_id, ep, flags, _stream()))
    return out


def gemm_multi(weights, X, spec, epilogues=None, ep0s=None, outs=None, ws=None, streamk=True, tile=None, pdl=False):
    """b200q_gemm_multi: up to four weights that read the same X (q | k | v [| h_to_4h with bias + GELU]) in ONE launch
    of the CTA-pair kernel.
    weights: [(packed, scale) | (packed, scale, zmin), ...]; epilogues / ep0s: per weight ("none" | "bias" | ..).
    Returns the outputs [dtype[..., M, N_i]]."""
    spec = _spec(spec)
    n = len(weights)
    epilogues = list(epilogues) if epilogues is not None else ["none"] * n
    ep0s = list(ep0s) if ep0s is not None else [None] * n
    K = X.shape[-1]
    M = X.numel() // K
    if X.dtype != spec.torch_dtype:
        raise ValueError(f"X must be {spec.dtype}")
    _need_cuda(X)
    arr = (_lib.GemmOp * n)()
    res, keep = [], []
    for i, w in enumerate(weights):
        packed, scale = w[0], w[1]
        zmin = w[2] if len(w) > 2 else None
        _need_cuda(packed, scale, zmin, ep0s[i])
        _check_min(spec, scale, zmin)
        N = _check_weight(spec, packed, scale, K)
        ep = _ep_id(epilogues[i])
        odt = torch.float32 if ep == EPILOGUES["cast_f32"] else spec.torch_dtype
        out = outs[i] if outs is not None else torch.empty(tuple(X.shape[:-1]) + (N,), dtype=odt, device=X.device)
        if out.numel() != M * N or out.dtype != odt:
            raise ValueError(f"out {i} must be {odt}[{M}, {N}]")
        arr[i] = _lib.GemmOp(packed.data_ptr(), scale.data_ptr(), _ptr(zmin), _ptr(ep0s[i]), None, None, out.data_ptr(), N, ep)
        res.append(out)
        keep.append((packed, scale, zmin))
    flags = GEMM_TILES[tile] | (FLAG_PDL if pdl else 0)
    flags |= FLAG_GEMM_STREAMK if streamk == "force" else (0 if streamk else FLAG_GEMM_NO_STREAMK)
    if ws is False:
        wp, wb = None, 0
    else:
        w_ = gemm_workspace(X.device) if ws is None else ws
        wp, wb = w_.data_ptr(), w_.numel() * w_.element_size()
    _lib.check(_lib.lib().b200q_gemm_multi(arr, n, X.data_ptr(), M, K, spec.fmt_id, flags, wp, wb, _stream()))
    return res


def nt_matmul(X, T, epilogue="none", ep0=None, ep1=None, ep2=None, out=None, ws=None, streamk=True, pdl=False):
    """NT_matmul: Y[m, n] = sum_k X[m, k] * T[n, k] (+ epilogue), f16, tcgen05 (ws / streamk / pdl as in gemm)."""
    ep = _ep_id(epilogue)
    _need_cuda(X, T, ep0, ep1, ep2, out)
    K = X.shape[-1]
    M = X.numel() // K
    N = T.shape[0]
    if T.shape[1] != K or X.dtype != torch.float16 or T.dtype != torch.float16:
        raise ValueError("nt_matmul: X f16[..., K], T f16[N, K]")
    odt = torch.float32 if ep == EPILOGUES["cast_f32"] else torch.float16
    if out is None:
        out = torch.empty(tuple(X.shape[:-1]) + (N,), dtype=odt, device=X.device)
    L = _lib.lib()
    flags = (FLAG_PDL if pdl else 0) | (FLAG_GEMM_STREAMK if streamk == "force" else (0 if streamk else FLAG_GEMM_NO_STREAMK))
    if ws is False:
        wp, wb = None, 0
    else:
        w_ = gemm_workspace(X.device) if ws is None else ws
        wp, wb = w_.data_ptr(), w_.numel() * w_.element_size()
    _lib.check(L.b200q_nt_matmul_ws(X.data_ptr(), T.data_ptr(), _ptr(ep0), _ptr(ep1), _ptr(ep2), out.data_ptr(),
                                    M, K, N, ep, flags, wp, wb, _stream()))
    return out


def launch_count() -> int:
    return int(_lib.lib().b200q_launch_count())


_WATCHDOG_STATUS = None


def program_watchdog(status: Optional[torch.Tensor] = None, timeout_ms: float = -1.0) -> None:
    """b200q_gemv_program_set_watchdog: bound every cross-CTA / cross-GPU wait of the program kernels.
    status: a pinned-host int64[1] tensor that receives 0xDEAD | site << 16 | step << 32 when a wait expires (the
    kernel then traps, so the launch fails instead of hanging); timeout_ms < 0: the default (10 s), 0: unbounded."""
    global _WATCHDOG_STATUS
    if status is not None and (status.dtype != torch.int64 or status.numel() < 1 or not status.is_pinned()):
        raise ValueError("status must be a pinned int64 tensor")
    _WATCHDOG_STATUS = status  # keep it alive
    _lib.lib().b200q_gemv_program_set_watchdog(None if status is None else status.data_ptr(),
                                               int(timeout_ms * 1e6) if timeout_ms >= 0 else -1)


class GemvProgram:
    """The GEMV chain of one decode step as ONE persistent launch (b200q_gemv_program_*).

    steps: list of dicts with keys
        weights   [(packed, scale), ...]   1..4 Linears that read the same activation
        x         dtype[K]                 activation (may be the `out` of an earlier step)
        x_mul     optional dtype[K]        activation = x * x_mul
        epilogues optional list            per Linear ("none", "add", ...)
        ep0s      optional list            per Linear epilogue operand or None
        outs      list                     per Linear output tensor (destination passing)
        reduces   optional list of bool    per Linear: row-parallel (tensor parallel; summed over the ranks in-program)
    tp = (rank, size, process_group): tensor-parallel program (b200q_gemv_program_build_tp / _run_tp): the
    workspace is allocated as symmetric memory so that the peers can push partial sums / finished values into it.
    Dependencies between steps are derived by the library from pointer overlap.  All tensors are kept
    alive by this object; run() launches on torch's current stream and is CUDA-graph capturable."""

    def __init__(self, steps, spec, tp=None):
        import ctypes

        import numpy as np

        spec = _spec(spec)
        self.spec = spec
        L = _lib.lib()
        n = len(steps)
        arr = (_lib.GemvStep * n)()
        tp_rank, tp_size, tp_group = tp if tp is not None else (0, 1, None)
        mask = np.zeros(n, dtype=np.uint8)
        self._keep = []
        dev = None
        for j, st in enumerate(steps):
            x = st["x"]
            x_mul = st.get("x_mul")
            weights = st["weights"]
            eps = list(st.get("epilogues") or ["none"] * len(weights))
            ep0s = list(st.get("ep0s") or [None] * len(weights))
            outs = st["outs"]
            for i, r in enumerate(st.get("reduces") or []):
                if r:
                    mask[j] |= 1 << i
            K = x.shape[-1]
            _need_cuda(x, x_mul)
            if x.numel() != K or x.dtype != spec.torch_dtype:
                raise ValueError(f"step {j}: x must be {spec.dtype}[K]")
            dev = x.device
            arr[j].n_ops = len(weights)
            arr[j].x = x.data_ptr()
            arr[j].x_mul = _ptr(x_mul)
            arr[j].K = K
            for i, (packed, scale) in enumerate(weights):
                _need_cuda(packed, scale, ep0s[i], outs[i])
                N = _check_weight(spec, packed, scale, K)
                ep = _ep_id(eps[i])
                odt = torch.float32 if (ep == EPILOGUES["cast_f32"] or spec.dtype == "float32") else torch.float16
                if outs[i].numel() != N or outs[i].dtype != odt:
                    raise ValueError(f"step {j}: out[{i}] must be {odt}[{N}]")
                arr[j].ops[i] = _lib.GemvOp(packed.data_ptr(), scale.data_ptr(), _ptr(ep0s[i]), outs[i].data_ptr(), N, ep)
                self._keep += [packed, scale, ep0s[i], outs[i]]
            self._keep += [x, x_mul]
        nbytes = int(L.b200q_gemv_program_bytes(n))
        self._host = np.zeros(nbytes // 8 + 1, dtype=np.uint64)  # 8-byte aligned host blob
        self._mask = mask
        _lib.check(L.b200q_gemv_program_build_tp(arr, n, spec.fmt_id, mask.ctypes.data, tp_rank, tp_size,
                                                 self._host.ctypes.data, nbytes))
        self._dev = torch.from_numpy(self._host.view(np.uint8)[:nbytes].copy()).to(dev)  # torch allocations are >= 256-byte aligned
        ws_bytes = int(L.b200q_gemv_program_workspace_bytes(self._host.ctypes.data))
        self._peers = None
        if tp_size > 1:
            import torch.distributed as dist
            import torch.distributed._symmetric_memory as symm_mem

            # symmetric allocation: the peers write partial sums and finished activations straight into it
            self._ws = symm_mem.empty(ws_bytes, dtype=torch.uint8, device=dev)
            self._ws.zero_()
            self._hdl = symm_mem.rendezvous(self._ws, group=tp_group if tp_group is not None else dist.group.WORLD)
            ptrs = [int(p) for p in self._hdl.buffer_ptrs]
            if ptrs[tp_rank] != self._ws.data_ptr():
                raise RuntimeError("symmetric memory: buffer_ptrs[rank] is not the local workspace")
            self._peers = (ctypes.c_void_p * tp_size)(*ptrs)
            torch.cuda.synchronize(dev)
            dist.barrier(group=tp_group)  # every rank's workspace is zeroed before anyone launches
        else:
            self._ws = torch.zeros(ws_bytes, dtype=torch.uint8, device=dev)
        self.n_steps = n

    def run(self):
        L = _lib.lib()
        _lib.check(L.b200q_gemv_program_run_tp(self._host.ctypes.data, self._dev.data_ptr(), self._ws.data_ptr(),
                                               self._ws.numel(), self._peers, _stream()))
