#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_program_parity_gpu.py tests/test_gemv_program_gpu.py tests/test_decode_stack_gpu.py -m gpu -x -q > gpurun_out/q8_tests.log 2>&1; echo "tests rc=$?"; tail -12 gpurun_out/q8_tests.log
timeout 300 python scripts/perf_probe.py --fmt q8f16_0 --model llama-13b --iters 30 --no-stream > gpurun_out/q8_probe.log 2>&1; grep PROBE gpurun_out/q8_probe.log || tail -5 gpurun_out/q8_probe.log
B200Q_PROG_TC8=0 timeout 300 python scripts/perf_probe.py --fmt q8f16_0 --model llama-13b --iters 10 --no-stream > gpurun_out/q8_probe_old.log 2>&1; grep PROBE gpurun_out/q8_probe_old.log || tail -5 gpurun_out/q8_probe_old.log
