#!/bin/bash
# round-2 artifacts for the shipped state: bench lines of every config, ncu launch list of the C4 step, full capture of both K2 launches
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for c in c4 c1 c2 c3 c5; do
  extra="--no-cpu-baseline"; [ $c = c4 ] && extra=""
  timeout 900 python bench.py --steps 20 --warmup 5 --config $c $extra > $O/r02y_bench_$c.json 2> $O/r02y_bench_$c.err; echo "bench $c rc=$?"
  python - $c <<'PY'
import json,sys
d=json.loads(open("gpurun_out/r02y_bench_%s.json"%sys.argv[1]).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("value %.3e ms %.4f e2e %.3e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "sum/step %.3f"%d["roofline"]["sum_over_step"], "frac %.3f"%d["roofline"]["frac"], "parity", d["parity"]["ok"], "launches", d["gpu_launches"])
PY
done
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 200 --csv --log-file $O/r02y_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02y_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:update_fused -s 12 -c 2 -f -o $O/r02y_k2 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02y_ncu_k2.log 2>&1; echo "ncu k2 rc=$?"
