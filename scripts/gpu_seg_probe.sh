#!/bin/bash
# fused / segmented token rate and the 1-step launch latency: main build against libb200q_$1.so
mkdir -p gpurun_out
for lib in "" $1; do
  if [ -n "$lib" ]; then export B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$lib.so; fi
  timeout 200 python scripts/perf_probe.py --iters 100 --no-stream 2>&1 | grep PROBE
  timeout 200 python scripts/perf_probe.py --iters 50 --no-stream --segments reference 2>&1 | grep PROBE
  timeout 200 python scripts/perf_probe.py --no-token --no-stream --big --big-n 16384 2>&1 | grep PROBE
done
if [ -n "$2" ]; then timeout 900 python -m pytest $2 -m gpu -x -q 2>&1 | tail -4; fi
