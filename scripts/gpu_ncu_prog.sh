#!/bin/bash
# ncu --set full of one decode-program launch (one token), after the same command ran clean without ncu
mkdir -p gpurun_out
timeout 300 python bench.py --steps 4 --warmup 3 --headline-only > gpurun_out/n_plain.json 2> gpurun_out/n_plain.err; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemv_program -s 6 -c 1 -o gpurun_out/r2_tc_prog_full -f python bench.py --steps 4 --warmup 3 --headline-only > gpurun_out/n_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out/r2_tc_prog_full.ncu-rep
