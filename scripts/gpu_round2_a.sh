#!/bin/bash
# GPU-box script (round 2, call A): tests, smoke, bench (both arms), then kernel diagnostics.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/a_smi.log 2>&1
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 ) > gpurun_out/a_tests.log
echo "tests: $(tail -1 gpurun_out/a_tests.log)"
( timeout 300 python __graft_entry__.py smoke 2>&1 | tail -5 ) > gpurun_out/a_smoke.log
echo "smoke: $(tail -1 gpurun_out/a_smoke.log)"
timeout 900 python bench.py --steps 100 --warmup 5 > gpurun_out/a_bench.json 2> gpurun_out/a_bench.err
echo "bench rc=$? $(head -c 300 gpurun_out/a_bench.json)"
timeout 400 python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/a_bench_ref.json 2> gpurun_out/a_bench_ref.err
echo "ref rc=$? $(head -c 200 gpurun_out/a_bench_ref.json)"
./scripts/ubench/tma_stream > gpurun_out/a_tma_stream.log 2>&1
echo "tma_stream rc=$?"; cat gpurun_out/a_tma_stream.log
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_timing.so timeout 300 python scripts/perf_probe.py --iters 3 > gpurun_out/a_timing.log 2>&1
echo "timing rc=$?"; grep -E "PROBE|^dq|^mma|^helper" gpurun_out/a_timing.log | sort | uniq -c | sort -rn | head -30
