timeout -s KILL 90 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-shape 2>/dev/null > gpurun_out/trace_raw.txt
grep "^TR" gpurun_out/trace_raw.txt | tail -n 1184 > gpurun_out/trace.txt
grep "^BT" gpurun_out/trace_raw.txt | tail -n 128 > gpurun_out/btrace.txt
wc -l gpurun_out/trace.txt gpurun_out/btrace.txt
