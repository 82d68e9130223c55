#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_rollout.py tests/test_gpu_kernels.py tests/test_gpu_dropin.py -m gpu -q -p no:cacheprovider > $O/r02i_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/r02i_pytest.log
