timeout -s KILL 60 python scripts/prog_sweep.py q4f16_0 4096x32000 2>&1 | tail -6
timeout -s KILL 60 python scripts/prog_sweep.py q4f16_0 4096x4096 2>&1 | tail -6
timeout -s KILL 90 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-per-shape 2>/dev/null | grep -E "^(dq|x|mma|helper):" | tail -5
