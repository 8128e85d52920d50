#!/bin/bash
# full single-GPU validation: tests, smoke, both bench arms
mkdir -p gpurun_out
T=${1:-x}
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 ) > gpurun_out/${T}_tests.log; echo "tests: $(tail -1 gpurun_out/${T}_tests.log)"
( timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5 ) > gpurun_out/${T}_smoke.log; echo "smoke: $(tail -1 gpurun_out/${T}_smoke.log)"
timeout 900 python bench.py > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"
timeout 400 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err; echo "ref rc=$?"
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/${T}_bench.json') if l.startswith('{')][-1])
print('HEADLINE', round(d['value'],1), 'tok/s frac', round(d['roofline']['frac'],4), 'e2e', round(d['e2e']['value'],1), 'parity', d['parity'] and (d['parity'].get('ok'), d['parity'].get('max_err_over_tol')))
print('segmented', d['segmented'] and (round(d['segmented']['value'],1), d['segmented']['logits_equal_to_headline']))
p=d['prefill']; print('prefill', p and (round(p.get('value',0),1), [ (s['shape'], s['tflops'], s.get('dequant_plus_cublas_tflops', s.get('three_launches_tflops'))) for s in p.get('per_shape',[])]))
for e in d['extra']: print('extra', json.dumps(e)[:300])
print('cpu', d.get('cpu_baseline',{}).get('value'))
PY
