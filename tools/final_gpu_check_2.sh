#!/bin/bash
# Two-GPU acceptance run: the multi-GPU test files, short tp2 / pp2 runs of the 6-layer test model through bench.py, and
# BERT-large dp2 fp16 with dynamic loss scaling + optimizer-state sharding (the losses of its last steps are in the
# JSON line).  Output in gpurun_out/.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_fused_tp_gpu.py tests/test_sdp_gpu.py -m gpu -x -q > gpurun_out/final_pytest_2gpu.log 2>&1
echo "pytest_exit=$?" >> gpurun_out/final_pytest_2gpu.log; tail -3 gpurun_out/final_pytest_2gpu.log
RUN_TIMEOUT=120 tools/run_n.sh 2 final_b2_tp2_test.log -- --model gptj-test --steps 3 --warmup 3 --parallelism tp2
RUN_TIMEOUT=120 tools/run_n.sh 2 final_b2_pp2_test.log SMP_SCHEDULE_RECORD_STEPS=4 -- --model gptj-test --steps 6 --warmup 3 --parallelism pp2
RUN_TIMEOUT=150 tools/run_n.sh 2 final_b2_bert_fp16.log -- --model bert-large --fp16 1 --shard-optimizer-state 1 --parallelism dp2 --steps 12 --warmup 3
