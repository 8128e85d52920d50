#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py tests/test_prefill_stack_gpu.py -m gpu -x -q > gpurun_out/gm_tests.log 2>&1; echo "tests rc=$?"; tail -8 gpurun_out/gm_tests.log
timeout 300 python scripts/gemm_multi_probe.py 2>&1 | grep MULTI
timeout 300 python scripts/gemm_gelu_probe.py 2>&1 | grep EP
