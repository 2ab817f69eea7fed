set -x
export SMP_LOG_LEVEL=warning
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/sdp_fused_check.py --small > gpurun_out/sdp_check_small.log 2>&1; echo "small rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 tools/sdp_fused_check.py > gpurun_out/sdp_check.log 2>&1; echo "big rc=$?"
timeout 900 python -m pytest tests/test_sdp_gpu.py -x -q -p no:logging > gpurun_out/sdp_test.log 2>&1; echo "test rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 2 --parallelism sdp2 > gpurun_out/bench_sdp2.log 2>&1; echo "bench sdp2 rc=$?"
tail -3 gpurun_out/sdp_check_small.log gpurun_out/sdp_check.log gpurun_out/sdp_test.log gpurun_out/bench_sdp2.log
