export SMP_LOG_LEVEL=warning
timeout 500 python bench.py --impl baseline --steps 3 --warmup 3 > gpurun_out/bench_n1_baseline.log 2>&1; echo "baseline rc=$?"; grep '"metric"\|unavailable\|\[baseline\]' gpurun_out/bench_n1_baseline.log | cut -c1-400
