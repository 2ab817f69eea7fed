#!/bin/bash
# Everything that was written after the last multi-GPU session and has run on CPU / gloo only (DESIGN.md section 4,
# "Verified on GPUs vs written afterwards").  Run on a box with 4 GPUs (gpurun --gpus 4); output lands in gpurun_out/.
#   1. the native communication runtime over NCCL (uneven all-to-all, p2p handles, bucket reducer)
#   2. the hybrid matrix (HF GPT-2 dense vs smp: tp4, tp2 x pp2, sdp4, optimizer-state sharding, offload ...)
#   3. the misaligned-operand GEMM test and the rest of the one-GPU tier
#   4. short tp2 / tp4 / pp2 x tp2 bench runs with the native reducer and, for comparison, with SMP_NATIVE_COMM=0
mkdir -p gpurun_out
export SMP_RUN_NATIVE_GPU=1 SMP_RUN_HYBRID_GPU=1
timeout 600 python -m pytest tests/test_native_torch_gloo.py -m gpu -x -q > gpurun_out/pending_native_runtime.log 2>&1
echo "native runtime: $(tail -1 gpurun_out/pending_native_runtime.log)"
timeout 1500 python -m pytest tests/test_hybrid.py -m gpu -q > gpurun_out/pending_hybrid.log 2>&1
echo "hybrid: $(tail -1 gpurun_out/pending_hybrid.log)"
timeout 900 python -m pytest tests -m gpu -x -q --deselect tests/test_hybrid.py --deselect tests/test_native_torch_gloo.py \
    > gpurun_out/pending_pytest_gpu.log 2>&1
echo "gpu tier: $(tail -1 gpurun_out/pending_pytest_gpu.log)"
RUN_TIMEOUT=300 tools/run_n.sh 2 pending_b2_tp2_native.log -- --steps 3 --warmup 3 --parallelism tp2
RUN_TIMEOUT=300 tools/run_n.sh 2 pending_b2_tp2_python.log SMP_NATIVE_COMM=0 -- --steps 3 --warmup 3 --parallelism tp2
RUN_TIMEOUT=300 tools/run_n.sh 4 pending_b4_tp4_native.log -- --steps 3 --warmup 3 --parallelism tp4
RUN_TIMEOUT=400 tools/run_n.sh 4 pending_b4_pp2tp2_native.log -- --steps 2 --warmup 3 --parallelism pp2tp2
