#!/bin/bash
mkdir -p gpurun_out
for v in "" _nowd _v1 _v2 _v3 _v4 _v5; do
  B200Q_LIB=$PWD/mlc_llm_b200/libb200q$v.so timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/e_probe$v.log 2>&1; echo "probe$v rc=$?"; grep PROBE gpurun_out/e_probe$v.log
done
