export SMP_LOG_LEVEL=warning
SMP_TEST_HANG_DUMP_S=110 SMP_TEST_TIMEOUT_S=140 timeout 600 python -m pytest tests -m gpu -x -q -p no:logging > gpurun_out/pytest_gpu_2gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/pytest_gpu_2gpu.log
