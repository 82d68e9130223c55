#!/bin/bash
# Round 2, GPU batch N (8 GPUs): exchange algorithms at N = 8 and 4; 2-GPU comm tests (both algorithms) on the same box.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > $O/r02n_bench_n1.json 2> $O/r02n_bench_n1.err; echo "bench n1 rc=$?"
for n in 8 4; do
 for a in owner push_all; do
  DPPO_COMM_ALGO=$a timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2972$n bench.py --gpus $n --steps 20 --warmup 5 > $O/r02n_bench_n${n}_$a.json 2> $O/r02n_bench_n${n}_$a.err; echo "bench n$n $a rc=$?"
 done
done
python - <<'PY'
import json
base=json.loads(open("gpurun_out/r02n_bench_n1.json").read().strip().splitlines()[-1])["value"]
print("N=1 value %.3e"%base)
for n in (8,4):
  for a in ("owner","push_all"):
    try:
        d=json.loads(open("gpurun_out/r02n_bench_n%d_%s.json"%(n,a)).read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
        print("N=%d %-8s value %.3e (x%.2f) ms %.4f e2e %.3e"%(n,a,d["value"], d["value"]/base, d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam","k4_welford")}, "parity", d["parity"]["ok"], d["parity"]["replicas_identical"], d["comm"]["algorithm"])
    except Exception as e: print(n, a, "failed", e)
PY
timeout 900 python -m pytest tests/test_gpu_comm.py -m gpu -q -p no:cacheprovider -k "two_gpus" > $O/r02n_pytest_2gpu.log 2>&1; echo "pytest 2gpu rc=$?"; tail -4 $O/r02n_pytest_2gpu.log
