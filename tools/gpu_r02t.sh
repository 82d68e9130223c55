#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_comm.py -m gpu -q -p no:cacheprovider -k "welford or shared_stats" > $O/r02t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02t_pytest.log
echo bulk-store; timeout 120 python tools/k4_probe.py | tail -2; timeout 120 python tools/k4_probe.py 65536 17 | tail -1
echo stg-store; DPPO_K4_STG=1 timeout 120 python tools/k4_probe.py | tail -2; DPPO_K4_STG=1 timeout 120 python tools/k4_probe.py 65536 17 | tail -1
