export SMP_LOG_LEVEL=warning
timeout 300 python -m pytest tests -m gpu -x -q -p no:logging > gpurun_out/pytest_gpu_1gpu.log 2>&1; echo "pytest rc=$?"; tail -n 2 gpurun_out/pytest_gpu_1gpu.log
timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/bench_n1.log | cut -c1-230
CMD="python bench.py --layers 2 --steps 1 --warmup 1"
timeout 150 $CMD > gpurun_out/plain_small.log 2>&1 && timeout 300 ncu --set full --clock-control none --import-source on -k regex:gemm_tcgen05 -s 60 -c 3 -o gpurun_out/prof_gemm_pair $CMD > gpurun_out/ncu_full.log 2>&1; echo "ncu rc=$?"
