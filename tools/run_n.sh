#!/bin/bash
# usage: tools/run_n.sh N LOGNAME [env VAR=... ] -- bench args      (one torchrun of bench.py on N GPUs of this box)
N=$1; LOG=$2; shift 2
ENVS=()
while [ "$1" != "--" ] && [ $# -gt 0 ]; do ENVS+=("$1"); shift; done
shift
PORT=$((29500 + RANDOM % 2000))
env "${ENVS[@]}" timeout ${RUN_TIMEOUT:-600} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N "$@" > gpurun_out/$LOG 2>&1
echo "== $LOG rc=$?"; grep -E '^\{' gpurun_out/$LOG | tail -1 | cut -c1-400; grep -E "Error|error:" gpurun_out/$LOG | head -3
