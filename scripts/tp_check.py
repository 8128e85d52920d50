"""torchrun --nproc-per-node N scripts/tp_check.py : fused in-program tensor-parallel reduction vs the NCCL chain."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from mlc_llm_b200.decode_stack import DecodeStack
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
layers = int(sys.argv[1]) if len(sys.argv) > 1 else 2
x = (torch.randn(4096, device=dev, generator=torch.Generator(device=dev).manual_seed(1)) * 0.05).half()
res = {}
for name, kw in (("nccl-chain", dict(use_program=False)), ("fused-program", dict(use_program=True))):
    st = DecodeStack("llama-7b", "q4f16_0", device=dev, tp_rank=rank, tp_size=world, layers=layers, **kw)
    st.x_in.copy_(x)
    st.capture()
    y = st.decode_step().clone()
    y2 = st.decode_step().clone()
    torch.cuda.synchronize()
    assert torch.equal(y, y2), f"{name}: replay differs"
    assert torch.isfinite(y).all()
    res[name] = y
    if rank == 0:
        print(name, "fused_tp" if st.fused_tp else "", "launches", st.launches_per_step, "collectives", st.collectives_per_step, "logits[:4]", y[:4].tolist(), flush=True)
    dist.barrier()
d = (res["nccl-chain"] - res["fused-program"]).abs().max().item()
m = res["nccl-chain"].abs().max().item()
if rank == 0:
    print(f"max |diff| {d:.4e}  max |logit| {m:.4e}  -> {'OK' if d <= 2e-2 * max(1.0, m) else 'MISMATCH'}", flush=True)
torch.cuda.synchronize()
os._exit(0)
