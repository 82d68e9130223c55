#!/bin/bash
# Round 2, GPU batch D: K4 after the fp32-recurrence rewrite (tests + timing), the release experiment under the replay
# stress (early1/early2 variants), full suite, bench c4.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -p no:cacheprovider > $O/r02d_pytest.log 2>&1; echo "pytest rc=$?"
tail -8 $O/r02d_pytest.log
timeout 300 python tools/k4_probe.py; timeout 300 python tools/k4_probe.py 65536 17; timeout 300 python tools/k4_probe.py 1048576 376
for v in early1 early2; do
  DPPO_LIB=tools/variants/libdppo_b200_$v.so timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -q -p no:cacheprovider -k "repeatable" > $O/r02d_stress_$v.log 2>&1; echo "stress $v rc=$?"; tail -3 $O/r02d_stress_$v.log
done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r02d_bench_c4.json 2> $O/r02d_bench_c4.err; echo "bench c4 rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02d_bench_c4.json")); k=d["roofline"]["kernels"]
print("ms", d["ms_per_step"], {n:k[n]["us"] for n in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, k["hbm_study"], d["parity"])
PY
