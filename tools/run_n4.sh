export SMP_LOG_LEVEL=warning
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 4 --steps 3 --warmup 3 > gpurun_out/bench_n4.log 2>&1; echo "tp4 rc=$?"
grep '"metric"' gpurun_out/bench_n4.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['config']['parallelism'], d['value'], d['ms_per_step'], d['model_tflops_per_gpu'], d['peak_mem_gb'], d['e2e']['value'])"
grep -v Warning gpurun_out/bench_n4.log | grep "Error" | sort | uniq -c | head -5
