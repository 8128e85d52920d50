"""Tensor-parallel decode program with the reduction fused into the launch (b200q_gemv_program_build_tp /
_run_tp): two ranks, peer pushes over NVLink, against the CPU oracle operator by operator (oracle/stack_check.py) and against the NCCL all-reduce chain on identical shards.
Needs >= 2 GPUs (skipped otherwise); the sharding arithmetic itself is covered on CPU by
test_tp_sharding_cpu.py (gloo, world_size 2)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_fused_tp_program_matches_nccl_chain():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "scripts", "tp_check.py"), "2"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-2000:]
    assert "fused-program fused_tp launches 1 collectives 1 parity True" in out, out[-2000:]
    assert "nccl-chain  launches" in out and "-> OK" in out, out[-2000:]
