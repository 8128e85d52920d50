timeout -s KILL 40 python scripts/dbg_tc.py 4096 4096 2>&1 | tail -1
timeout -s KILL 40 python scripts/dbg_tc.py 11008 4096 2>&1 | tail -1 || exit 1
timeout -s KILL 90 python -m pytest tests/test_gemv_program_gpu.py -x -q 2>&1 | tail -3
timeout -s KILL 60 python scripts/prog_sweep.py q4f16_0 2>&1 | tail -5
timeout -s KILL 90 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-per-shape 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('bench', d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'])"
