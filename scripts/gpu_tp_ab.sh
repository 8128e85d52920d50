#!/bin/bash
# usage: gpu_tp_ab.sh N suffix : headline-only TP bench, main build against libb200q_<suffix>.so
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541"
for lib in main $2; do
  if [ "$lib" != main ]; then export B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$lib.so; fi
  timeout 300 $TR bench.py --gpus $N --steps 100 --warmup 5 --headline-only > gpurun_out/tpab_$lib.json 2> gpurun_out/tpab_$lib.err; echo "$lib rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/tpab_$lib.json') if l.startswith('{')][-1])
    print('$lib TP$N', round(d['value'],1), 'tok/s', round(d['ms_per_step']*1e3,1), 'us')
except Exception as e:
    print('parse failed', e); print(open('gpurun_out/tpab_$lib.err').read()[-800:])
PY
done
if [ -n "$2" ]; then B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$2.so timeout 300 $TR scripts/tp_check.py 2 2>&1 | grep -E "fused-program|oracle parity"; fi
