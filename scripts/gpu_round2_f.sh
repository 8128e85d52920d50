#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gemm_gpu.py -x -q 2>&1 | tail -15 ) > gpurun_out/f_gemm_tests.log; cat gpurun_out/f_gemm_tests.log | tail -8
timeout 300 python scripts/gemm_probe.py > gpurun_out/f_gemm_probe.log 2>&1; echo "probe rc=$?"; grep GEMM gpurun_out/f_gemm_probe.log
for v in "" _nounroll; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q$v.so timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/f_probe$v.log 2>&1; grep PROBE gpurun_out/f_probe$v.log
done
