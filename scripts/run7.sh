B200Q_TC_DBG=32 timeout -s KILL 100 python scripts/prog_sweep.py q4f16_0 4096x32000 2>&1 | tail -5
