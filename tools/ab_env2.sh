#!/bin/bash
# Generic same-box A/B of one environment switch: per-op table (selected rows) and C2 solve time per value.
# bash tools/ab_env2.sh TAG VAR v1 v2 ...
TAG=$1; VAR=$2; shift 2; O=gpurun_out; mkdir -p $O
for v in "$@"; do
  env $VAR=$v timeout 200 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_$v.txt 20 > /dev/null 2>&1; echo "$VAR=$v: $(head -1 $O/${TAG}_ops_$v.txt)"
  grep -E "^(down0|down1)\.(qkv|rowmax|expscores|pv|attn_out|geglu|ff_out) " $O/${TAG}_ops_$v.txt | awk '{printf "%s %s  ", $1, $3} END{print ""}'
  env $VAR=$v timeout 200 python bench.py --core --steps 20 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('C2 $VAR=$v', round(d['ms_per_step'],3), 'ms')" | tee -a $O/${TAG}_ab.txt
done
