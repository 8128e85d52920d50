for i in 1 2; do
timeout -s KILL 60 python scripts/prog_sweep.py q4f16_0 4096x32000,4096x12288 2>&1 | tail -2
timeout -s KILL 90 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-per-shape 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"
done
