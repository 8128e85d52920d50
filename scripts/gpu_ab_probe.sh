#!/bin/bash
# A/B of builds of the library on the decode-program probe: scripts/gpu_ab_probe.sh "<suffix> <suffix> .." [pytest file]
# (mlc_llm_b200/libb200q_<suffix>.so, built with make BUILD=_build_<suffix> OUT=../libb200q_<suffix>.so EXTRA=..)
mkdir -p gpurun_out
timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/ab_main.log 2>&1
for sfx in $1; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$sfx.so timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/ab_$sfx.log 2>&1
done
grep -h PROBE gpurun_out/ab_main.log; for sfx in $1; do grep -h PROBE gpurun_out/ab_$sfx.log; done
timeout 200 python scripts/perf_probe.py --fmt q8f16_0 --iters 50 --no-stream 2>&1 | grep PROBE
if [ -n "$2" ]; then timeout 600 python -m pytest $2 -m gpu -x -q 2>&1 | tail -5; fi
