#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_comm.py -m gpu -q -p no:cacheprovider -k "welford or shared_stats" > $O/r02r_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 $O/r02r_pytest.log
timeout 120 python tools/k4_probe.py; timeout 120 python tools/k4_probe.py 16777216 4; timeout 120 python tools/k4_probe.py 65536 17 | tail -1
