#!/bin/bash
# GPU-box script: plain run, then the ncu launch list and one full capture of the GEMV kernel.
set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-shape"
$CMD > gpurun_out/ncu_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_l.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemv -s 300 -c 6 -f -o gpurun_out/prof_gemv $CMD > gpurun_out/ncu_f.log 2>&1
echo "full rc=$?"
tail -3 gpurun_out/ncu_f.log
ls -la gpurun_out
