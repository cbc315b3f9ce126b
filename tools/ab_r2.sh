#!/bin/bash
# Same-box A/B: round-1 library (libserenade_b200_r1.so, kept out of git) vs HEAD with its run-time switches.
# bash tools/ab_r2.sh TAG
TAG=${1:-ab}
O=gpurun_out; mkdir -p $O
run() {  # name, env...
  local name=$1; shift
  echo "=== $name ($*)"
  env "$@" timeout 300 python tests/diagnose.py --time 2>&1 | grep -E "B=.*graph=True"
  env "$@" timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_${name}.txt > /dev/null 2>&1; head -1 $O/${TAG}_ops_${name}.txt
}
[ -f serenade_b200/libserenade_b200_r1.so ] && run r1 SRN_LIB_VARIANT=r1
run head X=1
run ew0 SRN_EW16=0
run ew15 SRN_EW16=15
run nogroup SRN_GROUPED=0 SRN_EW16=0
[ -f serenade_b200/libserenade_b200_r1.so ] && run r1b SRN_LIB_VARIANT=r1
run head2 X=1
