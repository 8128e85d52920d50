timeout -s KILL 100 python scripts/dbg_tc.py 4096 4096 2>&1 | tail -1
timeout -s KILL 200 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-per-shape 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"
