timeout -s KILL 300 python -m pytest tests/test_gemv_program_gpu.py -x -q > gpurun_out/t_prog.log 2>&1; tail -12 gpurun_out/t_prog.log
