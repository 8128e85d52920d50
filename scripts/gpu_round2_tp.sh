#!/bin/bash
# usage: gpu_round2_tp.sh N
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
timeout 300 $TR scripts/tp_check.py 2 > gpurun_out/tp${N}_check.log 2>&1; echo "tp_check rc=$?"; grep -E "parity|diff|Error|error" gpurun_out/tp${N}_check.log | head
timeout 600 $TR bench.py --gpus $N --steps 100 --warmup 5 --no-prefill > gpurun_out/tp${N}_bench.json 2> gpurun_out/tp${N}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/tp${N}_bench.json') if l.startswith('{')][-1])
    print('TP${N}', round(d['value'],1), 'tok/s e2e', round(d['e2e']['value'],1), 'parity', d['parity'] and (d['parity']['ok'], d['parity']['max_err_over_tol']), 'launches', d['roofline']['launches_per_step'])
    for e in d['extra']: print('  extra', e['metric'], round(e['value'],1), e['config'].get('reduction'), round(e.get('frac_of_hbm_peak', e.get('frac_of_tensor_peak', 0)),3))
except Exception as e:
    print('bench parse failed', e); print(open('gpurun_out/tp${N}_bench.err').read()[-1500:])
PY
