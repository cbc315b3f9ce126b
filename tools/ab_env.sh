#!/bin/bash
# A/B on ONE box through an environment switch: bash tools/ab_env.sh TAG VAR=VALUE  (runs default, VAR=VALUE, default)
TAG=$1; KV=$2
O=gpurun_out; mkdir -p $O
for v in "" "$KV" ""; do
  name=${v:-base}
  echo "=== $name"
  env $v timeout 300 python tests/diagnose.py --time 2>&1 | grep -E "^v |B=" | grep -E "graph=True|^v "
  env $v timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_${name//=/_}.txt > /dev/null 2>&1; head -1 $O/${TAG}_ops_${name//=/_}.txt
done
