#!/bin/bash
# Round 2, GPU batch E: K2 with multi-tile tensor-core layers / trunk-serial tiles (C5), the exact round-1 racy kernel
# (f727d22) under the repeatability probes, bench c5 + c4.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests/test_gpu_kernels.py -m gpu -q --maxfail=40 -p no:cacheprovider -k "update" > $O/r02e_pytest_k2.log 2>&1; echo "pytest k2 rc=$?"
tail -15 $O/r02e_pytest_k2.log
for c in c5 c4; do
  timeout 600 python bench.py --steps 20 --warmup 5 --config $c --no-cpu-baseline > $O/r02e_bench_$c.json 2> $O/r02e_bench_$c.err; echo "bench $c rc=$?"; tail -3 $O/r02e_bench_$c.err
done
python - <<'PY'
import json
for c in ("c5","c4"):
    try:
        d=json.load(open("gpurun_out/r02e_bench_%s.json"%c)); k=d["roofline"]["kernels"]
        print(c,"ms", d["ms_per_step"], {n:k[n]["us"] for n in ("k2_update_fused","k3_reduce_adam","k1_gae_scan","k4_welford")}, "frac", d["roofline"]["frac"], d["parity"]["ok"], d["parity"]["theta_rel"], d["parity"]["loss_rel"])
    except Exception as e: print(c, "failed", e)
PY
V=tools/variants/libdppo_b200_f727d22.so
DPPO_LIB=$V timeout 300 python tools/stream_vs_steps.py > $O/r02e_race_f727d22.log 2>&1; echo "race f727d22 rc=$?"; cat $O/r02e_race_f727d22.log
DPPO_LIB=$V PROBE_REPS=200 timeout 600 python tools/repeat_probe.py > $O/r02e_probe_f727d22.log 2>&1; echo "probe f727d22 rc=$?"; grep "K2" $O/r02e_probe_f727d22.log
DPPO_LIB=$V timeout 600 python -m pytest tests/test_gpu_engine.py -m gpu -q -p no:cacheprovider -k "repeatable" > $O/r02e_stress_f727d22.log 2>&1; echo "stress f727d22 rc=$?"; tail -3 $O/r02e_stress_f727d22.log
