mkdir -p gpurun_out
N=${1:-2}
for prog in 1; do
timeout -s KILL 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --program $prog --no-cpu-baseline --no-per-shape > gpurun_out/bench_tp${N}_p${prog}.json 2> gpurun_out/bench_tp${N}_p${prog}.err
echo "rc=$?"; tail -1 gpurun_out/bench_tp${N}_p${prog}.json | python -c "
import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N', d['n_gpus'], 'prog', $prog, d['value'], d['ms_per_step'], d['roofline']['frac'], d['e2e']['value'], d['collectives_per_step'])" || tail -5 gpurun_out/bench_tp${N}_p${prog}.err
done
