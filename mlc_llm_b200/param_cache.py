"""The reference's on-disk parameter format: `params/ndarray-cache.json` + raw shards.

The reference quantizes the weights once on the host (`transform_params`, mlc_llm/utils.py:98-110), writes them with
`tvmjs.dump_ndarray_cache(param_dict, f"{artifact_path}/params", meta_data={"ParamSize": n}, encode_format="raw")`
(utils.py:113-128) and reads them back, positionally as param_0 .. param_{n-1}, in `load_params` (utils.py:131-139;
C++ side cpp/cli_main.cc:319-334, cpp/llm_chat.cc:826-861).  The format (TVM `python/tvm/contrib/tvmjs.py`):

    ndarray-cache.json = {"metadata": {"ParamSize": n, ...},
                          "records": [{"dataPath": "params_shard_0.bin", "format": "raw-shard", "nbytes": ..,
                                       "md5sum": "..",
                                       "records": [{"name": "param_0", "shape": [..], "dtype": "uint32",
                                                    "format": "raw" | "f32-to-bf16", "nbytes": .., "byteOffset": ..}, ..]},
                                      ..]}

A shard holds whole tensors back to back; a tensor of at least half the shard cap gets a shard of its own.
`load` returns the tensors in parameter order (CPU or straight onto a CUDA device); `dump` writes the same format
(so that artefacts produced here load in the reference and the other way round - tests/test_param_cache_cpu.py
checks both directions against TVM's own implementation).  `group_quantized` pairs the positional tensors back into
(packed, scale[, min]) triples - the order LiftTransformParams emits them in.
"""
from __future__ import annotations

import hashlib
import json
import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from .quant_spec import QuantSpec

CACHE_JSON = "ndarray-cache.json"  # (TVM renamed it tensor-cache.json later; both are read)
_NP_DTYPES = {"float32": np.float32, "float16": np.float16, "uint32": np.uint32, "int32": np.int32, "uint16": np.uint16,
              "int8": np.int8, "uint8": np.uint8, "int64": np.int64, "float64": np.float64, "bfloat16": np.uint16}


def _md5(path: str) -> str:
    h = hashlib.md5()
    with open(path, "rb") as f:
        for chunk in iter(lambda: f.read(1 << 20), b""):
            h.update(chunk)
    return h.hexdigest()


def _bf16_to_f32(u16: np.ndarray) -> np.ndarray:
    return (u16.astype(np.uint32) << np.uint32(16)).view(np.float32)


def dump(params: Sequence, cache_dir: str, meta: Optional[dict] = None, shard_cap_mb: int = 32, prefix: str = "params_shard") -> str:
    """save_params: tensors (torch or numpy) -> cache_dir/ndarray-cache.json + params_shard_*.bin, names param_i."""
    os.makedirs(cache_dir, exist_ok=True)
    cap = shard_cap_mb << 20
    shards, cur, cur_recs = [], bytearray(), []

    def commit(data, recs):
        path = f"{prefix}_{len(shards)}.bin"
        full = os.path.join(cache_dir, path)
        with open(full, "wb") as f:
            f.write(data)
        shards.append({"dataPath": path, "format": "raw-shard", "nbytes": len(data), "records": recs, "md5sum": _md5(full)})

    for i, t in enumerate(params):
        a = t.detach().cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)
        if a.dtype == np.int32 and isinstance(t, torch.Tensor):
            a = a.view(np.uint32)  # packed words / bf16 pairs travel as int32 views in torch; the format stores uint32
        data = np.ascontiguousarray(a).tobytes()
        rec = {"name": f"param_{i}", "shape": list(a.shape), "dtype": str(a.dtype), "format": "raw", "nbytes": len(data)}
        if len(cur) + len(data) >= cap:
            if len(data) * 2 >= cap:  # out-of-band: a shard of its own
                rec["byteOffset"] = 0
                commit(data, [rec])
                continue
            if cur:
                commit(bytes(cur), cur_recs)
                cur, cur_recs = bytearray(), []
        rec["byteOffset"] = len(cur)
        cur_recs.append(rec)
        cur += data
    if cur:
        commit(bytes(cur), cur_recs)
    md = {"ParamSize": len(params)}
    md.update(meta or {})
    path = os.path.join(cache_dir, CACHE_JSON)
    with open(path, "w") as f:
        json.dump({"metadata": md, "records": shards}, f, indent=1)
    return path


def load(cache_dir: str, device=None, verify_md5: bool = False) -> Tuple[List[torch.Tensor], dict]:
    """load_params: (param_0 .. param_{n-1} as torch tensors on `device`, metadata).  uint32 tensors (the packed words and
    the f32 scheme's bf16 scale|min pairs) come back as int32 views, the convention of mlc_llm_b200.ops."""
    path = cache_dir
    if not path.endswith(".json"):
        for name in (CACHE_JSON, "tensor-cache.json"):
            if os.path.exists(os.path.join(cache_dir, name)):
                path = os.path.join(cache_dir, name)
                break
        else:
            raise FileNotFoundError(f"no {CACHE_JSON} under {cache_dir}")
    base = os.path.dirname(path)
    info = json.load(open(path))
    by_name: Dict[str, torch.Tensor] = {}
    for shard in info["records"]:
        if shard.get("format") != "raw-shard":
            raise ValueError(f"shard format {shard.get('format')!r} (only raw-shard is defined)")
        full = os.path.join(base, shard["dataPath"])
        raw = open(full, "rb").read()
        if len(raw) != shard["nbytes"]:
            raise ValueError(f"{full}: {len(raw)} bytes on disk, {shard['nbytes']} in the index")
        if verify_md5 and "md5sum" in shard and _md5(full) != shard["md5sum"]:
            raise ValueError(f"{full}: md5 mismatch")
        for rec in shard["records"]:
            off, nb, dt, shape = rec["byteOffset"], rec["nbytes"], rec["dtype"], rec["shape"]
            if off + nb > len(raw):
                raise ValueError(f"{rec['name']}: record runs past the end of its shard")
            buf = raw[off:off + nb]
            if rec.get("format") == "f32-to-bf16" and dt == "float32":
                a = _bf16_to_f32(np.frombuffer(buf, dtype=np.uint16)).reshape(shape)
            else:
                if dt not in _NP_DTYPES:
                    raise ValueError(f"{rec['name']}: dtype {dt}")
                a = np.frombuffer(buf, dtype=_NP_DTYPES[dt]).reshape(shape)
            if a.dtype == np.uint32:
                a = a.view(np.int32)
            if a.dtype == np.uint16:
                a = a.view(np.int16)
            t = torch.from_numpy(a.copy())
            by_name[rec["name"]] = t.to(device) if device is not None else t
    n = int(info["metadata"]["ParamSize"])
    missing = [i for i in range(n) if f"param_{i}" not in by_name]
    if missing:
        raise ValueError(f"parameters missing from the cache: {missing[:8]}")
    return [by_name[f"param_{i}"] for i in range(n)], info["metadata"]


def group_quantized(params: Sequence[torch.Tensor], spec: QuantSpec):
    """Pair the positional tensors back into weights.  The quantization pass emits, per quantized parameter, the
    packed words followed by the scale (and, for the f16 asymmetric scheme, the min) - `emit_encoding`,
    mlc_llm/transform/quantization.py:332-360 - and leaves every other parameter (norm weights, biases) alone.
    Returns a list of ("linear", packed, scale, min | None, K, N) for the transposed (K/.., N) matmul layout,
    ("embedding", packed, scale, min | None, V, K) for the row-major take_decode layout, or ("plain", tensor)."""
    out, i = [], 0
    sdt = spec.scale_torch_dtype
    n_aux = 2 if spec.has_min else 1
    while i < len(params):
        p = params[i]
        aux = list(params[i + 1:i + 1 + n_aux])
        ok = (p.dtype == torch.int32 and p.dim() == 2 and len(aux) == n_aux and all(a.dim() == 2 and a.dtype == sdt for a in aux)
              and all(a.shape == aux[0].shape for a in aux))
        if ok:
            sc = aux[0]
            mn = aux[1] if spec.has_min else None
            wpg = spec.n_words(spec.group_size)  # words per group
            if p.shape[1] == sc.shape[1] and p.shape[0] == sc.shape[0] * wpg:  # [n_words, N] + [n_groups, N]
                out.append(("linear", p, sc, mn, sc.shape[0] * spec.group_size, p.shape[1]))
                i += 1 + n_aux
                continue
            if p.shape[0] == sc.shape[0] and p.shape[1] == sc.shape[1] * wpg:  # [V, n_words] + [V, n_groups]
                out.append(("embedding", p, sc, mn, p.shape[0], sc.shape[1] * spec.group_size))
                i += 1 + n_aux
                continue
        out.append(("plain", p))
        i += 1
    return out
