set -x
mkdir -p gpurun_out
cd scripts/ubench && nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o pipes pipes.cu && nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o unpack unpack.cu && cd ../..
./scripts/ubench/pipes > gpurun_out/ub_pipes.log 2>&1
./scripts/ubench/unpack > gpurun_out/ub_unpack.log 2>&1
timeout 300 python scripts/prog_sweep.py q4f16_0 > gpurun_out/prog_sweep.log 2>&1
timeout 300 python bench.py --steps 50 --warmup 5 --program 1 --no-cpu-baseline > gpurun_out/bench_prog.json 2> gpurun_out/bench_prog.err
timeout 300 python bench.py --steps 50 --warmup 5 --program 0 --no-cpu-baseline > gpurun_out/bench_chain.json 2> gpurun_out/bench_chain.err
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1
tail -3 gpurun_out/pytest_gpu.log
