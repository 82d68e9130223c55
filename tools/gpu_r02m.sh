#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_comm.py tests/test_gpu_kernels.py -m gpu -q -p no:cacheprovider -k "comm or loopback or owner or world_one or shared or welford" > $O/r02m_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 $O/r02m_pytest.log
timeout 300 python tools/k4_probe.py | tail -1; timeout 300 python tools/k4_probe.py 65536 17 | tail -1
