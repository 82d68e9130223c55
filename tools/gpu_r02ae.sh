#!/bin/bash
# one 8-GPU box: the scaling sweep N = 1, 2, 4, 8 the way the driver runs it
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > $O/r02ae_bench_n1.json 2> $O/r02ae_bench_n1.err; echo "bench n1 rc=$?"
for n in 2 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n bench.py --gpus $n --steps 20 --warmup 5 > $O/r02ae_bench_n$n.json 2> $O/r02ae_bench_n$n.err; echo "bench n$n rc=$?"
done
python - <<'PY'
import json
base=None
for n in (1,2,4,8):
    d=json.loads(open("gpurun_out/r02ae_bench_n%d.json"%n).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
    base=base or d["value"]
    print("N=%d value %.4e (x%.2f) ms %.4f e2e %.4e"%(n, d["value"], d["value"]/base, d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k4_welford")}, "parity", d["parity"]["ok"], d["parity"].get("replicas_identical"))
PY
