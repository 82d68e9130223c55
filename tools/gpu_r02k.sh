#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --maxfail=25 -p no:cacheprovider -k "welford or shared_stats or engine or rollout or dropin" > $O/r02k_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r02k_pytest.log
timeout 300 python tools/k4_probe.py; timeout 300 python tools/k4_probe.py 65536 17; timeout 300 python tools/k4_probe.py 1048576 376
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_fused -s 1 -c 1 -o $O/r02k_k4_big -f python tools/k4_probe.py > $O/r02k_ncu_k4.log 2>&1; echo "ncu k4 rc=$?"
