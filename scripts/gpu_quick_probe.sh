#!/bin/bash
mkdir -p gpurun_out
timeout 200 python scripts/perf_probe.py --iters 100 --big > gpurun_out/q_probe.log 2>&1; grep PROBE gpurun_out/q_probe.log
