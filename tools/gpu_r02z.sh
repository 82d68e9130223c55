#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -q -x -p no:cacheprovider -k "update or engine or cycle" > $O/r02z_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/r02z_pytest.log
for dwg in 1 0; do
  DPPO_K2_DWG=$dwg timeout 600 python bench.py --steps 20 --warmup 5 --config c5 --no-cpu-baseline > $O/r02z_bench_c5_dwg$dwg.json 2> $O/r02z_bench_c5_dwg$dwg.err; echo "bench c5 dwg=$dwg rc=$?"
  python - $dwg <<'PY'
import json,sys
d=json.loads(open("gpurun_out/r02z_bench_c5_dwg%s.json"%sys.argv[1]).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("value %.3e ms %.4f e2e %.3e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "sum/step %.3f"%d["roofline"]["sum_over_step"], "frac %.3f"%d["roofline"]["frac"], "parity", d["parity"]["ok"], "launches", d["gpu_launches"])
PY
done
