#!/bin/bash
# Two-GPU acceptance run: the multi-GPU test files, then short GPT-J 6B tp2 and pp2 bench runs.  Output in gpurun_out/.
mkdir -p gpurun_out
timeout 420 python -m pytest tests/test_fused_tp_gpu.py tests/test_sdp_gpu.py tests/test_offload_gpu.py -m gpu -x -q > gpurun_out/final_pytest_2gpu.log 2>&1
echo "pytest_exit=$?" >> gpurun_out/final_pytest_2gpu.log; tail -3 gpurun_out/final_pytest_2gpu.log
RUN_TIMEOUT=240 tools/run_n.sh 2 final_b2_tp2.log -- --steps 3 --warmup 3 --parallelism tp2
RUN_TIMEOUT=300 tools/run_n.sh 2 final_b2_pp2.log -- --steps 2 --warmup 3 --parallelism pp2
