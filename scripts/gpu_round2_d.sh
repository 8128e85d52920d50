#!/bin/bash
mkdir -p gpurun_out
for v in "" _nowd _nacc16; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q$v.so timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/d_probe$v.log 2>&1; echo "probe$v rc=$?"; grep PROBE gpurun_out/d_probe$v.log
done
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_trace.so timeout 200 python scripts/perf_probe.py --iters 1 --no-stream > gpurun_out/d_trace.log 2>&1; echo "trace rc=$?"
grep "^TR" gpurun_out/d_trace.log | tail -1184 > gpurun_out/d_trace_last.log
python scripts/trace_steps.py gpurun_out/d_trace_last.log | tee gpurun_out/d_trace_steps.txt
grep "^BT" gpurun_out/d_trace.log | tail -128 > gpurun_out/d_btrace_last.log
timeout 300 python scripts/diag_segments.py 32 > gpurun_out/d_diag32.log 2>&1; grep DIAG gpurun_out/d_diag32.log
