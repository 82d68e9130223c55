#!/bin/bash
# Round 2, GPU batch B: which kernel is not bit-repeatable (tools/repeat_probe.py), K1/K4 after the chain changes.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --maxfail=25 -p no:cacheprovider -k "welford or gae or noalloc or shared_stats or dropin" > $O/r02b_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 $O/r02b_pytest.log
timeout 900 python tools/repeat_probe.py > $O/r02b_probe.log 2>&1; echo "probe rc=$?"
cat $O/r02b_probe.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r02b_bench_c4.json 2> $O/r02b_bench_c4.err; echo "bench c4 rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02b_bench_c4.json")); k=d["roofline"]["kernels"]
print("ms", d["ms_per_step"], {n:k[n]["us"] for n in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, k["hbm_study"], d["parity"]["ok"])
PY
