#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/gemm_ncu_case.py > gpurun_out/h_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -c 8 -f -o gpurun_out/h_gemm python scripts/gemm_ncu_case.py > gpurun_out/h_ncu.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/h_ncu.log; ls -la gpurun_out/h_gemm.ncu-rep
