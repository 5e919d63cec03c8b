set -x
timeout 1200 python -m pytest tests -m gpu -v -k "multi_gpu or two_gpus" 2>&1 | grep -E "PASSED|FAILED|SKIPPED|passed|failed|Error" | sed 's/^.*:://' | tee gpurun_out/r2_pytest_multi_gpu_4gpu_final.txt
