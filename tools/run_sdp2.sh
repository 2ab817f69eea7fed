export SMP_LOG_LEVEL=warning
SMP_TEST_HANG_DUMP_S=100 SMP_TEST_TIMEOUT_S=130 timeout 420 python -m pytest tests/test_sdp_gpu.py -x -q -p no:logging > gpurun_out/sdp_test.log 2>&1; echo "test rc=$?"
tail -n 3 gpurun_out/sdp_test.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 3 --warmup 2 --parallelism sdp2 > gpurun_out/bench_sdp2.log 2>&1; echo "bench sdp2 rc=$?"
grep '"metric"' gpurun_out/bench_sdp2.log | cut -c1-900
SMP_SDP_FUSED=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 3 --warmup 2 --parallelism sdp2 > gpurun_out/bench_sdp2_nccl.log 2>&1; echo "bench sdp2 nccl rc=$?"
grep '"metric"' gpurun_out/bench_sdp2_nccl.log | cut -c1-900
