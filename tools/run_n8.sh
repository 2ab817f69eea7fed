export SMP_LOG_LEVEL=warning
run() { name=$1; shift; timeout 330 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus 8 --steps 3 --warmup 3 "$@" > gpurun_out/bench_n8_$name.log 2>&1; echo "$name rc=$?"; grep '"metric"' gpurun_out/bench_n8_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['parallelism'], d['value'], d['ms_per_step'], d['model_tflops_per_gpu'], d['peak_mem_gb'], d['e2e']['value'])" 2>/dev/null; }
PORT=29601 run pp2tp4
