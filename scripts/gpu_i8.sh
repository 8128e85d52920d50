#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_program_parity_gpu.py tests/test_gemv_program_gpu.py tests/test_decode_stack_gpu.py -m gpu -x -q > gpurun_out/i8_tests.log 2>&1; echo "tests rc=$?"; tail -25 gpurun_out/i8_tests.log
timeout 300 python scripts/perf_probe.py --iters 100 --big > gpurun_out/i8_probe.log 2>&1; grep PROBE gpurun_out/i8_probe.log || tail -5 gpurun_out/i8_probe.log
B200Q_PROG_I8=0 timeout 300 python scripts/perf_probe.py --iters 100 --big > gpurun_out/i8_probe_f16.log 2>&1; grep PROBE gpurun_out/i8_probe_f16.log || tail -5 gpurun_out/i8_probe_f16.log
