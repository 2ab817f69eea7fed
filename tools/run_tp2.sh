export SMP_LOG_LEVEL=warning
timeout 200 python -m pytest tests/test_offload_gpu.py -x -q -p no:logging > gpurun_out/offload_test.log 2>&1; echo "offload rc=$?"; tail -n 6 gpurun_out/offload_test.log
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29622 tools/sdp_fused_check.py > gpurun_out/sdp_check_pair.log 2>&1; echo "sdpcheck rc=$?"; grep -E "reduce|PASS|FAIL" gpurun_out/sdp_check_pair.log | tail -n 3
SMP_TEST_HANG_DUMP_S=100 SMP_TEST_TIMEOUT_S=130 timeout 300 python -m pytest tests/test_sdp_gpu.py -x -q -p no:logging > gpurun_out/sdp_test.log 2>&1; echo "sdp tests rc=$?"; tail -n 2 gpurun_out/sdp_test.log
