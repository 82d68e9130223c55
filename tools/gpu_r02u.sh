#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_dropin.py tests/test_gpu_kernels.py -m gpu -q -x -p no:cacheprovider > $O/r02u_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r02u_pytest.log
for pdl in 1 0 1 0; do
  DPPO_PDL=$pdl timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline > $O/r02u_bench_pdl$pdl.json 2> $O/r02u_bench_pdl$pdl.err; echo "bench pdl=$pdl rc=$?"
  python - <<PY
import json
d=json.loads(open("$O/r02u_bench_pdl$pdl.json").read().strip().splitlines()[-1])
print("pdl=$pdl value %.3fM ms/step %.4f e2e %.3fM parity %s k2 %s" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["parity"]["ok"], d["roofline"].get("kernel_us", d["roofline"].get("k2_us"))))
PY
done
