"""world_size-2 gloo test (CPU) of the tensor-parallel host logic: the column / row shards taken by
mlc_llm_b200.decode_stack.shard_packed, evaluated with the ORACLE on each rank and combined with one
all-reduce per sharded block, reproduce the unsharded oracle result."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, name, ret):
    sys.path.insert(0, ROOT)
    from mlc_llm_b200.decode_stack import shard_packed
    from mlc_llm_b200.quant_spec import SPECS
    from oracle import quant_oracle as qo

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fmt, spec = qo.FORMATS[name], SPECS[name]
        g = spec.group_size
        h, inter = 8 * g, 16 * g  # divisible by 2 * group
        rng = np.random.default_rng(7)  # same on both ranks
        def mk(K, N):
            W = (rng.standard_normal((N, K)) * 0.02).astype(fmt.np_dtype)
            return qo.encode_sym(W, fmt)
        w_up, w_down = mk(h, inter), mk(inter, h)
        x = (rng.standard_normal(h)).astype(fmt.np_dtype)
        res = (rng.standard_normal(h)).astype(fmt.np_dtype)
        # unsharded reference: a = x @ up ; y = res + a @ down   (f64 on the dequantized weights)
        up = qo.decode_sym(*w_up, h, fmt).astype(np.float64)
        down = qo.decode_sym(*w_down, inter, fmt).astype(np.float64)
        a_full = (x.astype(np.float64) @ up).astype(fmt.np_dtype)
        y_full = res.astype(np.float64) + a_full.astype(np.float64) @ down
        # rank-local: column shard of up, row shard of down
        as_t = lambda a: torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a)
        pu, su, Ku, Nu = shard_packed(as_t(w_up[0]), as_t(w_up[1]), spec, h, inter, col_slice=(rank, world))
        pd, sd, Kd, Nd = shard_packed(as_t(w_down[0]), as_t(w_down[1]), spec, inter, h, row_slice=(rank, world))
        assert (Ku, Nu, Kd, Nd) == (h, inter // world, inter // world, h)
        up_r = qo.decode_sym(pu.numpy().view(np.uint32), su.numpy(), Ku, fmt).astype(np.float64)
        down_r = qo.decode_sym(pd.numpy().view(np.uint32), sd.numpy(), Kd, fmt).astype(np.float64)
        # the shard is exactly the slice of the full dequantized matrix (no repacking happened)
        n0 = rank * Nu
        assert np.array_equal(up_r, up[:, n0:n0 + Nu]) and np.array_equal(down_r, down[n0:n0 + Kd, :])
        a_r = (x.astype(np.float64) @ up_r).astype(fmt.np_dtype)          # stays local between the two Linears
        part = a_r.astype(np.float64) @ down_r
        if rank == 0:
            part = part + res.astype(np.float64)                           # rank 0 carries the residual
        t = torch.from_numpy(part)
        dist.all_reduce(t)                                                 # ONE all-reduce per sharded block
        err = np.abs(t.numpy() - y_full).max()
        ret[rank] = float(err)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["q4f16_0", "q3f16_0", "q8f16_0"])
def test_column_row_sharding_with_allreduce_gloo(name):
    world = 2
    port = 29500 + (os.getpid() % 2000)
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, port, name, ret), nprocs=world, join=True)
        assert len(ret) == world
        assert max(ret.values()) < 1e-9, dict(ret)


def test_shard_rejects_partial_groups():
    sys.path.insert(0, ROOT)
    from mlc_llm_b200.decode_stack import shard_packed
    from mlc_llm_b200.quant_spec import SPECS

    spec = SPECS["q3f16_0"]
    p = torch.zeros((spec.n_words(4096), 64), dtype=torch.int32)
    s = torch.zeros((spec.n_groups(4096), 64), dtype=torch.float16)
    with pytest.raises(ValueError):
        shard_packed(p, s, spec, 4096, 64, row_slice=(0, 2))  # 2048 is not a multiple of 40
    with pytest.raises(ValueError):
        shard_packed(p, s, spec, 4096, 64, col_slice=(0, 16))  # 4 columns per shard
