export SMP_LOG_LEVEL=warning
timeout 150 python -m pytest tests/test_kernels_gpu.py -x -q -p no:logging -k "gemm or attention" > gpurun_out/gemm_test.log 2>&1; echo "gemm tests rc=$?"; tail -n 3 gpurun_out/gemm_test.log
timeout 200 python tools/gemm_shapes.py > gpurun_out/gemm_shapes_pair2.log 2>&1; echo "shapes rc=$?"; grep -E "QK|dP|PV|per-layer" gpurun_out/gemm_shapes_pair2.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.log 2>&1; echo "bench rc=$?"; grep '"metric"' gpurun_out/bench_n1.log | cut -c1-260
