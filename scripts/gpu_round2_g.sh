#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gemm_gpu.py -x -q 2>&1 | tail -5 ) > gpurun_out/g_gemm_tests.log; tail -2 gpurun_out/g_gemm_tests.log
timeout 300 python scripts/gemm_probe.py > gpurun_out/g_gemm_probe.log 2>&1; echo "probe rc=$?"; grep GEMM gpurun_out/g_gemm_probe.log
