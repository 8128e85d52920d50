#!/bin/bash
# Per-step event trace of the decode program (a -DB200Q_TC_TRACE build of the library, see scripts/README.md) and the
# consumer's ceiling: one GEMV whose weights stay resident in L2 (4096 -> 16384: 37 MB q4 / 70 MB q8), next to the
# HBM-bound one.  scripts/gpu_trace_probe.sh <suffix of the trace build>
mkdir -p gpurun_out
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_$1.so timeout 200 python scripts/perf_probe.py --iters 1 --no-stream > gpurun_out/trace_raw.log 2>&1
grep -c '^TR' gpurun_out/trace_raw.log
timeout 60 python scripts/trace_steps.py gpurun_out/trace_raw.log > gpurun_out/trace_steps.txt 2>&1
grep '^TR' gpurun_out/trace_raw.log | tail -n 1184 > gpurun_out/trace_tr_last_launch.txt
cat gpurun_out/trace_steps.txt
for f in q4f16_0 q8f16_0; do
  timeout 120 python scripts/perf_probe.py --fmt $f --no-token --no-stream --big --big-n 16384 2>&1 | grep PROBE
  timeout 120 python scripts/perf_probe.py --fmt $f --no-token --no-stream --big 2>&1 | grep PROBE
done
