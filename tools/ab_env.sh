#!/bin/bash
# Same-box A/B of run-time switches: bash tools/ab_env.sh TAG "NAME=ENV=VAL,ENV2=VAL" ...   (NAME=X=1 for the default)
TAG=$1; shift
O=gpurun_out; mkdir -p $O
for spec in "$@" "$1"; do
  name=${spec%%=*}; envs=${spec#*=}
  echo "=== $name ($envs)"
  env ${envs//,/ } timeout 300 python tests/diagnose.py --time 2>&1 | grep -E "B=.*graph=True"
  env ${envs//,/ } timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_${name}.txt > /dev/null 2>&1; head -1 $O/${TAG}_ops_${name}.txt
done
