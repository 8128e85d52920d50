#!/bin/bash
mkdir -p gpurun_out
for v in "" _nowd; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q$v.so timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/c_probe$v.log 2>&1; echo "probe$v rc=$?"; grep PROBE gpurun_out/c_probe$v.log
done
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_trace.so timeout 200 python scripts/perf_probe.py --iters 1 --no-stream > gpurun_out/c_trace.log 2>&1; echo "trace rc=$?"
grep -c "^TR" gpurun_out/c_trace.log
# the last launch's lines only (one launch = 148 CTAs x 8 steps)
grep "^TR" gpurun_out/c_trace.log | tail -1184 > gpurun_out/c_trace_last.log
python scripts/trace_steps.py gpurun_out/c_trace_last.log | tee gpurun_out/c_trace_steps.txt
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_timing.so timeout 200 python scripts/perf_probe.py --iters 2 --no-token --no-stream --big > gpurun_out/c_timing_big.log 2>&1
grep -E "PROBE|^dq|^mma|^helper" gpurun_out/c_timing_big.log | sort | uniq -c | sort -rn | head -8
timeout 300 python scripts/diag_segments.py 32 > gpurun_out/c_diag32.log 2>&1; grep DIAG gpurun_out/c_diag32.log
