export SMP_LOG_LEVEL=warning
timeout 200 python -m pytest tests/test_kernels_gpu.py -x -q -p no:logging -k "fused_mlp" > gpurun_out/mlp_test.log 2>&1; echo "mlp test rc=$?"; tail -n 8 gpurun_out/mlp_test.log
timeout 200 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/smoke.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/bench_n1.log | cut -c1-260
