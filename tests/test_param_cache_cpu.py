"""The on-disk parameter format (mlc_llm/utils.py:113-139 save_params / load_params = tvmjs ndarray cache): our
writer's artefacts load in TVM's own tvmjs loader and TVM's writer's artefacts load here, byte for byte; shard
splitting, out-of-band tensors, bf16-encoded f32, positional pairing back into (packed, scale[, min])."""
import json
import os
import shutil

import numpy as np
import pytest
import torch

from mlc_llm_b200 import param_cache
from mlc_llm_b200.quant_spec import SPECS
from oracle import quant_oracle as qo


def _params(rng):
    """A miniature transformed-parameter list in the reference's order: embedding (row-major), two Linears, a norm."""
    fmt = qo.FORMATS["q4f16_0"]
    out = []
    emb = qo.synth_weight(50, 64, fmt, seed=1)
    p, s = qo.encode_sym_rowmajor(emb, fmt)
    out += [p, s]
    for i, (N, K) in enumerate([(64, 64), (96, 64)]):
        p, s = qo.encode_sym(qo.synth_weight(N, K, fmt, seed=2 + i), fmt)
        out += [p, s]
    out.append(rng.standard_normal(64).astype(np.float16))  # a norm weight: not quantized
    big = rng.integers(0, 2**32, size=(520, 1024), dtype=np.uint64).astype(np.uint32)  # 2.1 MB: out-of-band at a 2 MB cap
    out.append(big)
    return out


def test_round_trip_and_sharding(tmp_path):
    rng = np.random.default_rng(0)
    params = _params(rng)
    d = str(tmp_path / "params")
    param_cache.dump(params, d, meta={"model": "tiny"}, shard_cap_mb=2)
    info = json.load(open(os.path.join(d, "ndarray-cache.json")))
    assert info["metadata"]["ParamSize"] == len(params) and info["metadata"]["model"] == "tiny"
    assert len(info["records"]) >= 2 and all(s["format"] == "raw-shard" for s in info["records"])
    assert any(len(s["records"]) == 1 and s["records"][0]["byteOffset"] == 0 and s["records"][0]["name"] == f"param_{len(params) - 1}"
               for s in info["records"])  # the big tensor sits alone in its shard
    back, meta = param_cache.load(d, verify_md5=True)
    assert meta["ParamSize"] == len(params)
    for a, t in zip(params, back):
        b = t.numpy()
        if a.dtype == np.uint32:
            b = b.view(np.uint32)
        assert a.shape == b.shape and a.tobytes() == b.tobytes()
    recs = param_cache.group_quantized(back, SPECS["q4f16_0"])
    assert [r[0] for r in recs] == ["embedding", "linear", "linear", "plain", "plain"]
    assert recs[0][4:] == (50, 64) and recs[1][4:] == (64, 64) and recs[2][4:] == (64, 96)


def test_asym_f16_triples_pair_up(tmp_path):
    fmt = qo.ASYM_FORMATS["q4f16_asym"]
    p, s, m = qo.encode_asym(qo.synth_weight(64, 128, fmt, seed=1), fmt)
    d = str(tmp_path / "params")
    param_cache.dump([p, s, m, np.ones(4, np.float16)], d)
    recs = param_cache.group_quantized(param_cache.load(d)[0], SPECS["q4f16_asym"])
    assert [r[0] for r in recs] == ["linear", "plain"] and recs[0][3] is not None and recs[0][4:] == (128, 64)


def test_corrupt_cache_is_rejected(tmp_path):
    d = str(tmp_path / "params")
    param_cache.dump([np.arange(10, dtype=np.float32)], d)
    with open(os.path.join(d, "params_shard_0.bin"), "ab") as f:
        f.write(b"x")
    with pytest.raises(ValueError):
        param_cache.load(d)


def _tvmjs():
    try:
        import tilelang  # noqa: F401  (puts its bundled tvm on sys.path)
        import tvm
        from tvm.contrib import tvmjs
        return tvm, tvmjs
    except Exception:  # noqa: BLE001
        return None, None


def test_interoperates_with_tvm_own_implementation(tmp_path):
    """Both directions against python/tvm/contrib/tvmjs.py of the TVM bundled in this image."""
    tvm, tvmjs = _tvmjs()
    if tvmjs is None:
        pytest.skip("no TVM in this environment")
    rng = np.random.default_rng(1)
    params = _params(rng)
    # ours -> TVM's loader
    d = str(tmp_path / "ours")
    param_cache.dump(params, d, shard_cap_mb=2)
    shutil.copy(os.path.join(d, "ndarray-cache.json"), os.path.join(d, "tensor-cache.json"))  # the newer TVM's file name
    loaded, meta = tvmjs.load_tensor_cache(d, tvm.cpu())
    assert int(meta["ParamSize"]) == len(params)
    for i, a in enumerate(params):
        assert loaded[f"param_{i}"].numpy().tobytes() == a.tobytes()
    # TVM's writer -> ours (incl. the f32-to-bf16 encoding)
    d2 = str(tmp_path / "theirs")
    f32 = rng.standard_normal((8, 16)).astype(np.float32)
    tvmjs.dump_tensor_cache({f"param_{i}": a for i, a in enumerate(params)}, d2, meta_data={"ParamSize": len(params)},
                            encode_format="raw", shard_cap_mb=2)
    back, _ = param_cache.load(d2)
    for a, t in zip(params, back):
        b = t.numpy()
        assert (b.view(np.uint32) if a.dtype == np.uint32 else b).tobytes() == a.tobytes()
    d3 = str(tmp_path / "bf16")
    tvmjs.dump_tensor_cache({"param_0": f32}, d3, meta_data={"ParamSize": 1}, encode_format="f32-to-bf16")
    (t,), _ = param_cache.load(d3)
    ref, _ = tvmjs.load_tensor_cache(d3, tvm.cpu())
    assert t.numpy().tobytes() == ref["param_0"].numpy().tobytes()
