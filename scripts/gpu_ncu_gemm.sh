#!/bin/bash
mkdir -p gpurun_out
timeout 300 python scripts/gemm_ncu_case.py > gpurun_out/g_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc2 -s 8 -c 4 -o gpurun_out/r2_gemm_streamk -f python scripts/gemm_ncu_case.py > gpurun_out/g_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r2_gemm_streamk.ncu-rep
