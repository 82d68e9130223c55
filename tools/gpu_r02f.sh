#!/bin/bash
# Round 2, GPU batch F: the round-1 racy kernel with its split results pinned before the early release.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for v in f727d22_pin f727d22; do
  V=tools/variants/libdppo_b200_$v.so
  DPPO_LIB=$V PROBE_REPS=300 timeout 600 python tools/repeat_probe.py > $O/r02f_probe_$v.log 2>&1; echo "probe $v rc=$?"; grep "K2 tensor" $O/r02f_probe_$v.log
  DPPO_LIB=$V timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -q -p no:cacheprovider -k "repeatable" > $O/r02f_stress_$v.log 2>&1; echo "stress $v rc=$?"; tail -4 $O/r02f_stress_$v.log
done
