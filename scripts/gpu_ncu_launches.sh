#!/bin/bash
# launch list (per-kernel durations) of the default bench command's headline part
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --headline-only > gpurun_out/l_plain.json 2> gpurun_out/l_plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gemv_program|gemv_stream|gemm_tc" -c 60 --csv --log-file gpurun_out/r2_launches.csv python bench.py --steps 2 --warmup 3 --headline-only > gpurun_out/l_ncu.log 2>&1; echo "ncu rc=$?"
wc -l gpurun_out/r2_launches.csv
