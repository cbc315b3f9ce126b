#!/bin/bash
# Why is one estimator call inside the solve slower than the sum of its isolated op timings?  The same per-op table with
# 5 launches per op (bursts at the maximum clock) and with 300 (each op sustained for tens of milliseconds, like the solve),
# SM clock and power sampled beside both; then the 128-column-tile A/B (SRN_BN128_COST).
# bash tools/gap_probe.sh TAG
TAG=${1:-gap}; O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=timestamp,clocks.sm,power.draw,clocks_throttle_reasons.active --format=csv -lms 200 > $O/${TAG}_smi.csv 2>&1 &
SMI=$!
for reps in 5 300; do
  date +%T.%N >> $O/${TAG}_marks.txt
  timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_reps$reps.txt $reps > /dev/null 2>&1; echo "reps $reps rc=$?"; head -1 $O/${TAG}_ops_reps$reps.txt
done
date +%T.%N >> $O/${TAG}_marks.txt
timeout 300 python bench.py --core --steps 20 --warmup 5 > $O/${TAG}_core.json 2>/dev/null; echo "core rc=$?"; cat $O/${TAG}_core.json | head -c 600; echo
date +%T.%N >> $O/${TAG}_marks.txt
kill $SMI
for cost in 0.45; do
  SRN_BN128_COST=$cost timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_bn128_$cost.txt 5 > /dev/null 2>&1; echo "bn128 $cost rc=$?"; head -1 $O/${TAG}_ops_bn128_$cost.txt
  SRN_BN128_COST=$cost timeout 300 python bench.py --core --steps 20 --warmup 5 > $O/${TAG}_core_bn128_$cost.json 2>/dev/null; head -c 400 $O/${TAG}_core_bn128_$cost.json; echo
done
