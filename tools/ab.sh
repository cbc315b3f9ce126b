#!/bin/bash
# A/B on ONE box: default library vs experiment variants (SRN_LIB_VARIANT).  bash tools/ab.sh TAG variant [variant...]
TAG=$1; shift
O=gpurun_out; mkdir -p $O
for v in "" "$@" ""; do
  name=${v:-base}
  echo "=== $name"
  SRN_LIB_VARIANT=$v timeout 300 python tests/diagnose.py --time 2>&1 | grep -E "^v |B=" | grep -E "graph=True|^v "
  SRN_LIB_VARIANT=$v timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_${name}.txt > /dev/null 2>&1; head -1 $O/${TAG}_ops_${name}.txt
done
