"""torchrun --nproc-per-node N scripts/tp_check.py [layers]: the tensor-parallel decode step on N GPUs, both routes
(the in-program one-hop reduction of the tcgen05 program; one NCCL all-reduce per sharded block), each checked
OPERATOR BY OPERATOR against the CPU oracle on rank 0 (oracle/stack_check.py: shards gathered, full weights
regenerated from the same seeds), plus replay determinism and the two routes against each other."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from mlc_llm_b200.decode_stack import DecodeStack
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
layers = int(sys.argv[1]) if len(sys.argv) > 1 else 2
model = sys.argv[2] if len(sys.argv) > 2 else "llama-7b"
x = (torch.randn(8192, device=dev, generator=torch.Generator(device=dev).manual_seed(1)) * 0.05).half()
res, ok = {}, True
full = None
if rank == 0:
    from oracle import quant_oracle as qo, stack_check
    full = DecodeStack(model, "q4f16_0", device=dev, layers=layers, use_program=False)
for name, kw in (("nccl-chain", dict(use_program=False)), ("fused-program", dict(use_program=True))):
    st = DecodeStack(model, "q4f16_0", device=dev, tp_rank=rank, tp_size=world, layers=layers, **kw)
    st.x_in.copy_(x[: st.shape.hidden])
    st.capture()
    y = st.decode_step().clone()
    y2 = st.decode_step().clone()
    torch.cuda.synchronize()
    assert torch.equal(y, y2), f"{name}: replay differs"
    assert torch.isfinite(y).all()
    res[name] = y
    snap = st.snapshot()
    if rank == 0:
        def weights(li, nm):
            w = full.lm_head if nm == "lm_head" else full.layers[li][nm]
            return w.packed.cpu().numpy().view("uint32"), w.scale.cpu().numpy()
        rec = stack_check.check_launches(snap, weights, qo.FORMATS["q4f16_0"], group_scaled=bool(kw["use_program"]))
        # the NCCL chain rounds every rank's partial sum to f16 before the all-reduce (the reduction runs on f16
        # tensors): up to ~tp x the single-rounding bound, reported but not required; the in-program reduction adds
        # the f32 partials and must meet the bound
        if kw["use_program"]:
            ok = ok and rec["ok"]
        print(name, "fused_tp" if st.fused_tp else "", "launches", st.launches_per_step, "collectives", st.collectives_per_step,
              "parity", rec["ok"], rec["max_err_over_tol"], rec["worst_op"], "ops", rec["ops_checked"], flush=True)
    dist.barrier()
    del st
d = (res["nccl-chain"] - res["fused-program"]).abs().max().item()
m = res["nccl-chain"].abs().max().item()
if rank == 0:
    good = ok and d <= 2e-2 * max(1.0, m)
    print(f"max |diff| {d:.4e}  max |logit| {m:.4e}  oracle parity {ok} -> {'OK' if good else 'MISMATCH'}", flush=True)
torch.cuda.synchronize()
os._exit(0)
