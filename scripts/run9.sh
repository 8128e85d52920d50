timeout -s KILL 90 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-per-shape 2>/dev/null > gpurun_out/trace_raw.txt
grep -c "^TR" gpurun_out/trace_raw.txt
tail -n 2400 gpurun_out/trace_raw.txt | grep "^TR" | tail -n 1184 > gpurun_out/trace.txt
wc -l gpurun_out/trace.txt
