mkdir -p gpurun_out
timeout -s KILL 240 python -m pytest tests/test_gemv_program_gpu.py -x -q -k "q4f16_0 or graph" > gpurun_out/t_prog.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_prog.log
tail -25 gpurun_out/t_prog.log
timeout -s KILL 200 python scripts/prog_sweep.py q4f16_0 > gpurun_out/prog_sweep_tc.log 2>&1; echo "sweep rc=$?" >> gpurun_out/prog_sweep_tc.log
cat gpurun_out/prog_sweep_tc.log
