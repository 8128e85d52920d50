mkdir -p gpurun_out
B="python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-per-shape"
timeout -s KILL 120 $B > gpurun_out/b_plain.log 2>&1 || exit 1
timeout -s KILL 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemv_program -c 40 --csv --log-file gpurun_out/launches_prog.csv $B > gpurun_out/ncu_ll.log 2>&1
timeout -s KILL 600 ncu --set full --clock-control none --import-source on -k regex:gemv_program -s 6 -c 1 -f -o gpurun_out/tc_prog_full $B > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/launches_prog.csv | cut -c1-60
