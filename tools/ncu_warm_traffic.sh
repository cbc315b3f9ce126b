#!/bin/bash
# DRAM traffic of a C2 solve with the caches left alone (ncu --cache-control none): what each kernel really reads from and
# writes to HBM when its inputs were produced by the previous kernel, beside the cold-cache list of r02_launches_c2_bench.txt.
# bash tools/ncu_warm_traffic.sh TAG
TAG=${1:-warm}; O=gpurun_out; mkdir -p $O
timeout 120 python bench.py --core --steps 1 --warmup 3 > $O/${TAG}_plain.json 2>/dev/null; echo "plain rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --cache-control none --clock-control none -c 2600 --csv \
    --log-file $O/${TAG}_launches.csv python bench.py --core --steps 1 --warmup 3 > $O/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
python tools/launch_summary.py $O/${TAG}_launches.csv $O/${TAG}_launches.txt | head -24
grep -m3 -i "pass" $O/${TAG}_ncu.log
