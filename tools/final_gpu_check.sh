#!/bin/bash
# One-GPU acceptance run: the GPU test tier, the train-vs-fp32-twin loss check, the flagship bench, smoke(), then ONE
# ncu --set full capture of the attention kernels (forward, dQ, dK/dV; each after its plain run has exited 0).
# Everything lands in gpurun_out/.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/final_pytest_gpu.log 2>&1; echo "pytest_exit=$?" >> gpurun_out/final_pytest_gpu.log
tail -3 gpurun_out/final_pytest_gpu.log
timeout 300 python tools/loss_check.py --steps 6 > gpurun_out/final_loss_check.log 2>&1; echo "exit=$?" >> gpurun_out/final_loss_check.log
grep -E "OK|MISMATCH|exit=|Error" gpurun_out/final_loss_check.log | tail -9
timeout 400 python bench.py --steps 5 --warmup 3 > gpurun_out/final_bench_n1.log 2>&1; echo "== bench rc=$?"
grep -E '^\{' gpurun_out/final_bench_n1.log | tail -1 | cut -c1-600
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; tail -1 gpurun_out/final_smoke.log
timeout 200 python tools/attn_check.py --bwd --time > gpurun_out/final_attn_check.log 2>&1; echo "attn_check rc=$?"
grep -E "^time|^clocks|ALL OK|FAIL" gpurun_out/final_attn_check.log | head -30
if timeout 120 python tools/attn_prof.py gptj > gpurun_out/final_attn_prof_plain.log 2>&1; then
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:attn_ --launch-skip 5 -c 5 -o gpurun_out/final_attn --force-overwrite \
      python tools/attn_prof.py gptj > gpurun_out/final_attn_ncu.log 2>&1
  echo "ncu rc=$?"
fi
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv,noheader > gpurun_out/final_clocks.txt 2>&1
