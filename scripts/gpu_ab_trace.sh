#!/bin/bash
# A/B of library builds on the decode-program probe, then the per-step trace of a -DB200Q_TC_TRACE build and a pytest
# file: scripts/gpu_ab_trace.sh "<suffix> .." <trace suffix> [pytest file]
mkdir -p gpurun_out
timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/ab_main.log 2>&1
for sfx in $1; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$sfx.so timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/ab_$sfx.log 2>&1
done
grep -h PROBE gpurun_out/ab_main.log; for sfx in $1; do grep -h PROBE gpurun_out/ab_$sfx.log; done
timeout 200 python scripts/perf_probe.py --iters 100 --no-stream --segments reference 2>&1 | grep PROBE
timeout 200 python scripts/perf_probe.py --fmt q8f16_0 --model llama-13b --iters 50 --no-stream 2>&1 | grep PROBE
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$2.so timeout 200 python scripts/perf_probe.py --iters 1 --no-stream > gpurun_out/trace_raw.log 2>&1
grep '^TR' gpurun_out/trace_raw.log | tail -n 1184 > gpurun_out/trace_tr_last_launch_new.txt
timeout 60 python scripts/trace_steps.py gpurun_out/trace_tr_last_launch_new.txt > gpurun_out/trace_steps_new.txt 2>&1
cat gpurun_out/trace_steps_new.txt
if [ -n "$3" ]; then timeout 600 python -m pytest $3 -m gpu -x -q 2>&1 | tail -5; fi
