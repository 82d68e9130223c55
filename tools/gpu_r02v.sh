#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_kernels.py -m gpu -q -x -p no:cacheprovider -k "update" > $O/r02v_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $O/r02v_pytest.log
