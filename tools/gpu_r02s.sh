#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 120 python tools/k4_probe.py | tail -1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:wf_fused -s 1 -c 1 -f -o $O/r02s_k4 python tools/k4_probe.py > $O/r02s_ncu.log 2>&1; echo "ncu rc=$?"
