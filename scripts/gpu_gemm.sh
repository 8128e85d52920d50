#!/bin/bash
# GEMM tests + probe
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -m gpu -x -q > gpurun_out/gemm_tests.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/gemm_tests.log
timeout 600 python scripts/gemm_probe.py > gpurun_out/gemm_probe.log 2>&1; echo "probe rc=$?"; grep GEMM gpurun_out/gemm_probe.log
