for d in 0 1 3 11; do echo "B200Q_TC_DBG=$d"; B200Q_TC_DBG=$d timeout -s KILL 60 python scripts/prog_sweep.py q4f16_0 4096x32000,4096x4096 2>&1 | tail -2; done
