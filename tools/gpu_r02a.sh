#!/bin/bash
# Round 2, GPU batch A: full GPU test suite, benches of every config, reference arm, release-race experiment,
# sanitizer passes, ncu launch list + full captures.  Everything lands in gpurun_out/.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/r02a_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -p no:cacheprovider > $O/r02a_pytest.log 2>&1; echo "pytest rc=$?" >> $O/r02a_pytest.log
tail -5 $O/r02a_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/r02a_bench_c4.json 2> $O/r02a_bench_c4.err; echo "bench c4 rc=$?"
for c in c1 c2 c3 c5; do
  timeout 600 python bench.py --steps 20 --warmup 5 --config $c --no-cpu-baseline > $O/r02a_bench_$c.json 2> $O/r02a_bench_$c.err; echo "bench $c rc=$?"
done
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02a_ref.json 2> $O/r02a_ref.err; echo "ref rc=$?"
for v in ship early1 early2; do
  if [ $v = ship ]; then L=""; else L="tools/variants/libdppo_b200_$v.so"; fi
  DPPO_LIB=$L timeout 300 python tools/stream_vs_steps.py > $O/r02a_race_$v.log 2>&1; echo "race $v rc=$?"
done
for t in racecheck synccheck; do
  timeout 600 compute-sanitizer --tool $t python tools/k2_sanitize.py 96 > $O/r02a_san_$t.log 2>&1; echo "sanitizer $t rc=$?"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 150 -c 160 --csv --log-file $O/r02a_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02a_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_fused -s 12 -c 1 -o $O/r02a_k2 -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02a_ncu_k2.log 2>&1; echo "ncu k2 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wf_fused -s 3 -c 1 -o $O/r02a_k4 -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02a_ncu_k4.log 2>&1; echo "ncu k4 rc=$?"
ls -la $O | tail -30
