#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r02o_bench_c4.json 2> $O/r02o_bench_c4.err; echo "bench c4 rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02o_bench_c4.json").read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("value %.3e ms %.4f e2e %.3e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "sum/step %.3f"%d["roofline"]["sum_over_step"], "parity", d["parity"]["ok"])
PY
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py -m gpu -q -p no:cacheprovider > $O/r02o_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r02o_pytest.log
