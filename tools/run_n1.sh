export SMP_LOG_LEVEL=warning
timeout 200 python -m pytest tests/test_kernels_gpu.py -x -q -p no:logging -k "attention or gemm" > gpurun_out/attn_test.log 2>&1; echo "attn tests rc=$?"; tail -n 12 gpurun_out/attn_test.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/bench_n1.log | cut -c1-260
