timeout -s KILL 100 python scripts/dbg_tc.py 4096 4096 2>&1 | tail -2
timeout -s KILL 100 python scripts/dbg_tc.py 11008 4096 2>&1 | tail -1
timeout -s KILL 100 python scripts/prog_sweep.py q4f16_0 2>&1 | tail -5
