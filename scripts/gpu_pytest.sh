#!/bin/bash
# usage: gpu_pytest.sh <tag> <pytest args...>
mkdir -p gpurun_out
T=$1; shift
( timeout 900 python -m pytest "$@" -x -q 2>&1 | tail -40 ) > gpurun_out/${T}_pytest.log; tail -25 gpurun_out/${T}_pytest.log
