#!/bin/bash
# two GPUs: the multi-GPU tests and the N = 2 bench line with the two-launch update
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_comm.py tests/test_gpu_dropin.py -m gpu -q -p no:cacheprovider > $O/r02x_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $O/r02x_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 > $O/r02x_bench_n2.json 2> $O/r02x_bench_n2.err; echo "bench n2 rc=$?"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02x_bench_n2.json").read().strip().splitlines()[-1]); k=d["roofline"]["kernels"]
print("N=2 value %.3e ms %.4f e2e %.3e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), {m:round(k[m]["us"],1) for m in ("k2_update_fused","k3_reduce_adam")}, "parity", d["parity"]["ok"], d["parity"].get("replicas_identical"), d.get("comm"))
PY
