#!/bin/bash
# 8-GPU box: N = 8 and N = 4 bench lines with the shipped kernels
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
for n in 8 4; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n bench.py --gpus $n --steps 20 --warmup 5 > $O/r02aa_bench_n$n.json 2> $O/r02aa_bench_n$n.err; echo "bench n$n rc=$?"
python - $n <<'PY'
import json,sys
d=json.loads(open("gpurun_out/r02aa_bench_n%s.json"%sys.argv[1]).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("N=%s value %.3e ms %.4f e2e %.3e"%(sys.argv[1], d["value"], d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k4_welford")}, "parity", d["parity"]["ok"], d["parity"].get("replicas_identical"), d.get("comm",{}).get("algorithm"))
PY
done
