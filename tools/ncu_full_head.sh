#!/bin/bash
# ncu --set full captures of the GEMM and row kernels of the second (warm) eager estimator call at C2, from the library in
# the tree; summaries via tools/ncu_summary.py.  bash tools/ncu_full_head.sh TAG
TAG=${1:-h}; O=gpurun_out; mkdir -p $O
timeout 120 python tools/one_forward.py 16 1000 2 > $O/${TAG}_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_kernel --launch-skip 60 --launch-count 32 -f \
    -o $O/${TAG}_gemm_full python tools/one_forward.py 16 1000 2 > $O/${TAG}_ncu_gemm.log 2>&1; echo "ncu-gemm rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gn_|layernorm|pack_cf' --launch-skip 21 --launch-count 12 -f \
    -o $O/${TAG}_row_full python tools/one_forward.py 16 1000 2 > $O/${TAG}_ncu_row.log 2>&1; echo "ncu-row rc=$?"
for r in gemm row; do
  ncu -i $O/${TAG}_${r}_full.ncu-rep --page raw --csv > $O/${TAG}_${r}_full_raw.csv 2>/dev/null
  python tools/ncu_summary.py $O/${TAG}_${r}_full_raw.csv $O/${TAG}_${r}_summary.txt | cut -c1-170
done
rm -f $O/${TAG}_row_full.ncu-rep $O/${TAG}_gemm_full.ncu-rep
