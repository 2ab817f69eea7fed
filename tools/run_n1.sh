export SMP_LOG_LEVEL=warning
timeout 600 python -m pytest tests -m gpu -x -q -p no:logging > gpurun_out/pytest_gpu_1gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu_1gpu.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/bench_n1.log | cut -c1-250
timeout 400 python bench.py --impl baseline --steps 3 --warmup 3 > gpurun_out/bench_n1_baseline.log 2>&1; echo "baseline rc=$?"; grep '"metric"' gpurun_out/bench_n1_baseline.log | cut -c1-250
CMD="python bench.py --layers 2 --steps 1 --warmup 1"
timeout 200 $CMD > gpurun_out/plain_small.log 2>&1 && timeout 500 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 40 -c 3 -o gpurun_out/prof_gemm $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"; ls -la gpurun_out/prof_gemm.ncu-rep
