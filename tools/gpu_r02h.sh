#!/bin/bash
# Round 2, GPU batch H: device sampler tests, e2e with device vs host sampler.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_sampler.py tests/test_gpu_engine.py -m gpu -q -p no:cacheprovider > $O/r02h_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r02h_pytest.log
for c in c4 c2; do
 for m in "" "--host-sampler"; do
  timeout 600 python bench.py --steps 20 --warmup 5 --config $c --no-cpu-baseline $m > $O/r02h_bench_$c$m.json 2> $O/r02h_bench_$c$m.err; echo "bench $c $m rc=$?"; tail -2 $O/r02h_bench_$c$m.err
  python - "$O/r02h_bench_$c$m.json" <<'PY'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("  value %.3e ms %.4f | e2e %.3e ms %.4f h2d %d | parity %s"%(d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_bytes_per_step"], d["parity"]["ok"]))
PY
 done
done
