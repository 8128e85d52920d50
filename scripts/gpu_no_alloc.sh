#!/bin/bash
# the no-allocation entry point's test: alone in a fresh process, then behind the whole GPU suite
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/na_alone.log 2>&1; echo "alone rc=$?"; tail -4 gpurun_out/na_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/na_all.log 2>&1; echo "all rc=$?"; tail -6 gpurun_out/na_all.log
