#!/bin/bash
# Round 2, GPU batch L (8 GPUs): weak scaling of C4 with the parity object at N = 8 and 4 (and 1 for the same box).
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi -L | wc -l
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r02l_bench_n1.json 2> $O/r02l_bench_n1.err; echo "bench n1 rc=$?"
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2971$n bench.py --gpus $n --steps 20 --warmup 5 > $O/r02l_bench_n$n.json 2> $O/r02l_bench_n$n.err; echo "bench n$n rc=$?"; tail -2 $O/r02l_bench_n$n.err
done
python - <<'PY'
import json
base=None
for n in (1,2,4,8):
    try:
        d=json.loads(open("gpurun_out/r02l_bench_n%d.json"%n).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
        if n==1: base=d["value"]
        print("N=%d value %.3e (x%.2f) ms %.4f e2e %.3e"%(n,d["value"], d["value"]/base, d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "parity", d["parity"]["ok"], d["parity"]["replicas_identical"], "%.1e"%d["parity"]["theta_rel"])
    except Exception as e: print(n, "failed", e)
PY
