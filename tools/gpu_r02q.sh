#!/bin/bash
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_engine.py tests/test_gpu_rollout.py -m gpu -q -p no:cacheprovider > $O/r02q_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r02q_pytest.log
for c in c4 c1; do
timeout 600 python bench.py --steps 20 --warmup 5 --config $c --no-cpu-baseline > $O/r02q_bench_$c.json 2> $O/r02q_bench_$c.err; echo "bench $c rc=$?"
python - $c <<'PY'
import json,sys
d=json.loads(open("gpurun_out/r02q_bench_%s.json"%sys.argv[1]).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("value %.3e ms %.4f e2e %.3e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "sum/step %.3f"%d["roofline"]["sum_over_step"], "parity", d["parity"]["ok"])
PY
done
