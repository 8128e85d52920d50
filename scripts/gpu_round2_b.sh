#!/bin/bash
# GPU-box script (round 2, call B): diagnostics. EVERY command under its own timeout.
mkdir -p gpurun_out
timeout 60 ./scripts/ubench/tma_stream > gpurun_out/b_tma_stream.log 2>&1; echo "tma_stream rc=$?"; cat gpurun_out/b_tma_stream.log
timeout 200 python scripts/diag_segments.py 4 > gpurun_out/b_diag_segments.log 2>&1; echo "diag rc=$?"; grep DIAG gpurun_out/b_diag_segments.log
for v in "" _nowd; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q$v.so timeout 200 python scripts/perf_probe.py --iters 100 > gpurun_out/b_probe$v.log 2>&1; echo "probe$v rc=$?"; grep PROBE gpurun_out/b_probe$v.log
done
B200Q_LIB=$PWD/mlc_llm_b200/libb200q.so timeout 200 python scripts/perf_probe.py --iters 100 --no-stream > gpurun_out/b_probe_again.log 2>&1; grep PROBE gpurun_out/b_probe_again.log
B200Q_LIB=$PWD/mlc_llm_b200/libb200q_timing.so timeout 200 python scripts/perf_probe.py --iters 3 > gpurun_out/b_timing.log 2>&1
echo "timing rc=$?"; grep -E "PROBE|^dq|^mma|^helper" gpurun_out/b_timing.log | sort | uniq -c | sort -rn | head -40
# compute-sanitizer on small cases (2-layer program, stream kernel, GEMM)
export B200Q_WATCHDOG_MS=0
timeout 500 compute-sanitizer --tool memcheck --log-file gpurun_out/b_memcheck.log python scripts/sanitize_cases.py > gpurun_out/b_memcheck.out 2>&1; echo "memcheck rc=$?"; tail -3 gpurun_out/b_memcheck.log
timeout 500 compute-sanitizer --tool synccheck --log-file gpurun_out/b_synccheck.log python scripts/sanitize_cases.py > gpurun_out/b_synccheck.out 2>&1; echo "synccheck rc=$?"; tail -3 gpurun_out/b_synccheck.log
timeout 500 compute-sanitizer --tool racecheck --log-file gpurun_out/b_racecheck.log python scripts/sanitize_cases.py > gpurun_out/b_racecheck.out 2>&1; echo "racecheck rc=$?"; tail -3 gpurun_out/b_racecheck.log
