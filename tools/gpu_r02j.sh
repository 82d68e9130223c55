#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
DPPO_LIB=tools/variants/libdppo_b200_timing.so timeout 300 python tools/k2clock.py > $O/r02j_k2clock.log 2>&1; echo "k2clock rc=$?"; cat $O/r02j_k2clock.log
timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -p no:cacheprovider > $O/r02j_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r02j_pytest.log
cp $O/closeness_report.json $O/r02j_closeness_full.json
