#!/bin/bash
# One GPU-box pass: parity tests, bench, per-op table, ncu launch list, ncu full captures (exported to CSV).
# Outputs in gpurun_out/ (kept under the 64 MiB pull limit: big .ncu-rep files are dropped after export).
set -u
TAG=${1:-r01}
WHAT=${2:-all}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest.log
tail -5 $O/${TAG}_pytest.log
timeout 300 python tests/diagnose.py --tf32 > $O/${TAG}_diag_tf32.log 2>&1; echo "diag tf32 rc=$?"; cp $O/diagnose.txt $O/${TAG}_diagnose_tf32.txt
grep -E "final|lengths" $O/${TAG}_diagnose_tf32.txt | head
timeout 300 python tests/diagnose.py --time > $O/${TAG}_diag_bf16.log 2>&1; echo "diag bf16 rc=$?"; cp $O/diagnose.txt $O/${TAG}_diagnose_bf16.txt
grep "B=" $O/${TAG}_diagnose_bf16.txt
SRN_DISABLE_PDL=1 timeout 300 python tests/diagnose.py --time > $O/${TAG}_diag_nopdl.log 2>&1; echo "diag nopdl rc=$?"
grep "B=" $O/${TAG}_diag_nopdl.log
timeout 600 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -c 2500 $O/${TAG}_bench.json
timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_c2.txt > /dev/null 2>&1; echo "ops rc=$?"
if [ "$WHAT" = "all" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 2000 --csv --log-file $O/${TAG}_launches.csv \
    python bench.py --steps 1 --warmup 3 > $O/${TAG}_ncu_launch.log 2>&1; echo "ncu-launch rc=$?"
# full captures of the second (warm) eager estimator call: 30 GEMM launches, 14 row-kernel launches
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel --launch-skip 73 --launch-count 30 -f \
    -o $O/${TAG}_gemm_full python tools/one_forward.py 16 1000 2 > $O/${TAG}_ncu_gemm.log 2>&1; echo "ncu-gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gn_|layernorm|pack_cf' --launch-skip 21 --launch-count 12 -f \
    -o $O/${TAG}_row_full python tools/one_forward.py 16 1000 2 > $O/${TAG}_ncu_row.log 2>&1; echo "ncu-row rc=$?"
for r in gemm row; do
  ncu -i $O/${TAG}_${r}_full.ncu-rep --page raw --csv > $O/${TAG}_${r}_full_raw.csv 2>/dev/null
done
fi
du -sm $O
# stay under the pull limit
if [ $(du -sm $O | cut -f1) -gt 60 ]; then rm -f $O/${TAG}_row_full.ncu-rep; fi
if [ $(du -sm $O | cut -f1) -gt 60 ]; then rm -f $O/${TAG}_gemm_full.ncu-rep; fi
ls -la $O
