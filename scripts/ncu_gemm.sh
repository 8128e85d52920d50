#!/bin/bash
set -x
CMD="python scripts/gemm_bench.py --shapes 2048x4096x11008 --fmts q4f16_0 --cublas 1"
$CMD > gpurun_out/ncu_gemm_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/ncu_gemm_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 2 -c 1 -f -o gpurun_out/prof_gemm_dense $CMD > gpurun_out/ncu_g1.log 2>&1
echo "rc=$?"
ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 48 -c 1 -f -o gpurun_out/prof_gemm_q4 $CMD > gpurun_out/ncu_g2.log 2>&1
echo "rc=$?"
ncu --set full --clock-control none -k regex:nvjet -s 2 -c 1 -f -o gpurun_out/prof_gemm_cublas $CMD > gpurun_out/ncu_g3.log 2>&1
echo "rc=$?"
tail -3 gpurun_out/ncu_g1.log gpurun_out/ncu_g2.log gpurun_out/ncu_g3.log
