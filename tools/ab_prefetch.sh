#!/bin/bash
# A/B of the L2 operand prefetch (SRN_PREFETCH_DIST = K blocks ahead; 0 = off) on one box: per-op table and C2 / C5 solves.
# The switch existed only in the experiment build measured in profiles/r02_ab_prefetch.txt (the prefetch was slower and was
# removed: DESIGN.md 3.4); the script is kept as the record of how those numbers were taken.
TAG=${1:-pf}; shift; O=gpurun_out; mkdir -p $O
for d in "$@"; do
  SRN_PREFETCH_DIST=$d timeout 200 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_pf$d.txt 20 > /dev/null 2>&1; head -1 $O/${TAG}_ops_pf$d.txt
  grep -E "^(down0|down1)\.(qkv|expscores|pv|attn_out|geglu|ff_out) " $O/${TAG}_ops_pf$d.txt | awk '{printf "%s %s  ", $1, $3} END{print ""}'
  SRN_PREFETCH_DIST=$d timeout 200 python bench.py --core --steps 20 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('C2 pf=$d', round(d['ms_per_step'],3), 'ms')" | tee -a $O/${TAG}_ab.txt
done
for d in "$@"; do
  SRN_PREFETCH_DIST=$d timeout 300 python bench.py --config C5 --core --steps 3 --warmup 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('C5 pf=$d', round(d['ms_per_step'],2), 'ms')" | tee -a $O/${TAG}_ab.txt
done
