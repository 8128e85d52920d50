N=${1:-2}
timeout -s KILL 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 scripts/tp_check.py 2 2>&1 | grep -v "^W\|^\*\*\*" | tail -12
