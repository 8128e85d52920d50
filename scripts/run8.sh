# ncu launch list of the bench command (program kernel launches only) -> gpurun_out/
timeout -s KILL 120 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-per-shape > gpurun_out/b_plain.log 2>&1 || exit 1
timeout -s KILL 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:gemv_program -c 40 --csv --log-file gpurun_out/launches_prog.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-per-shape > gpurun_out/ncu_ll.log 2>&1
tail -3 gpurun_out/launches_prog.csv
tail -1 gpurun_out/b_plain.log | cut -c1-400
