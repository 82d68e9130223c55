#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x -p no:cacheprovider > $O/r02w_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $O/r02w_pytest.log
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
for dwg in 1 0; do
  DPPO_K2_DWG=$dwg timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/r02w_bench_dwg$dwg.json 2> $O/r02w_bench_dwg$dwg.err; echo "bench dwg=$dwg rc=$?"
  python - <<PY
import json
d=json.loads(open("$O/r02w_bench_dwg$dwg.json").read().strip().splitlines()[-1])
r=d["roofline"]
print("dwg=$dwg value %.3fM ms/step %.4f e2e %.3fM parity %s launches %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["parity"], d["gpu_launches"]))
print({k:v for k,v in r.items() if k.endswith("_us") or k in ("sum_over_step","frac")})
PY
done
