#!/bin/bash
# Round 2, GPU batch P: the evidence set of the round - full GPU suite, smoke, benches of every config, reference arm,
# ncu launch list + full captures of K2 and K4, per-phase clocks of K2.
set -u
cd "$(dirname "$0")/.."
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q --maxfail=25 -p no:cacheprovider > $O/r02p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/r02p_pytest.log
cp $O/closeness_report.json $O/r02p_closeness.json
timeout 300 python __graft_entry__.py smoke > $O/r02p_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $O/r02p_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > $O/r02p_bench_c4.json 2> $O/r02p_bench_c4.err; echo "bench c4 rc=$?"
for c in c1 c2 c3 c5; do
  timeout 600 python bench.py --steps 20 --warmup 5 --config $c --no-cpu-baseline > $O/r02p_bench_$c.json 2> $O/r02p_bench_$c.err; echo "bench $c rc=$?"
done
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > $O/r02p_ref.json 2> $O/r02p_ref.err; echo "ref rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 --config c5 > $O/r02p_ref_c5.json 2> $O/r02p_ref_c5.err; echo "ref c5 rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 160 --csv --log-file $O/r02p_launches.csv \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02p_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_fused -s 12 -c 1 -o $O/r02p_k2 -f \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-parity > $O/r02p_ncu_k2.log 2>&1; echo "ncu k2 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:update_fused -s 12 -c 1 -o $O/r02p_k2_c5 -f \
  python bench.py --steps 2 --warmup 3 --config c5 --no-cpu-baseline --no-graph --no-parity > $O/r02p_ncu_k2_c5.log 2>&1; echo "ncu k2 c5 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:wf_fused -s 1 -c 1 -o $O/r02p_k4 -f python tools/k4_probe.py > $O/r02p_ncu_k4.log 2>&1; echo "ncu k4 rc=$?"
DPPO_LIB=tools/variants/libdppo_b200_timing.so timeout 300 python tools/k2clock.py > $O/r02p_k2clock.log 2>&1; echo "k2clock rc=$?"
ls -la $O | grep r02p
