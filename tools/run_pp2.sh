export SMP_LOG_LEVEL=warning
for W in 4 6; do
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2951$W bench.py --gpus 2 --steps 3 --warmup 2 --parallelism pp2 --active-microbatches $W > gpurun_out/bench_pp2_w$W.log 2>&1; echo "pp2 w$W rc=$?"
grep '"metric"' gpurun_out/bench_pp2_w$W.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['model_tflops_per_gpu'], d['peak_mem_gb'])"
done
