#!/bin/bash
# Round 2, GPU batch G (2 GPUs): the real one-process-per-GPU exchange path: tests + bench with the parity object.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_comm.py -m gpu -q -p no:cacheprovider > $O/r02g_pytest_comm.log 2>&1; echo "pytest comm rc=$?"; tail -15 $O/r02g_pytest_comm.log
for c in c4 c2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 20 --warmup 5 --config $c > $O/r02g_bench_n2_$c.json 2> $O/r02g_bench_n2_$c.err; echo "bench n2 $c rc=$?"; tail -3 $O/r02g_bench_n2_$c.err
done
python - <<'PY'
import json
for c in ("c4","c2"):
    try:
        d=json.load(open("gpurun_out/r02g_bench_n2_%s.json"%c)); k=d["roofline"]["kernels"]
        print(c,"value %.3e ms %.4f e2e %.3e"%(d["value"], d["ms_per_step"], d["e2e"]["value"]), {n:round(k[n]["us"],1) for n in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, d["parity"])
    except Exception as e: print(c, "failed", e)
PY
