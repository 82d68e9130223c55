#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_comm.py tests/test_gpu_rollout.py -m gpu -q -p no:cacheprovider -k "welford or shared_stats or rollout or obs" > $O/r02ad_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/r02ad_pytest.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r02ad_bench.json 2>/dev/null; echo rc=$?
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02ad_bench.json").read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("value %.3e ms %.4f"%(d["value"], d["ms_per_step"]), {m:round(k[m]["us"],2) for m in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "parity", d["parity"]["ok"], d["parity"]["obs_rel"], k["hbm_study"]["k4_welford_T2^22xO17"])
PY
