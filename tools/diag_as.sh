O=gpurun_out; mkdir -p $O
run() { name=$1; shift; env "$@" timeout 300 python tools/profile_ops.py 16 1000 $O/d_$name.txt > /dev/null 2>&1; echo "$name: $(head -1 $O/d_$name.txt)"; }
run as_full SRN_X=1
run as_noepi SRN_DBG_MODE=4
run plain_full SRN_ASTAT=0
run plain_noepi SRN_ASTAT=0 SRN_DBG_MODE=4
run onebox_plain SRN_ASTAT=0 SRN_LIB_VARIANT=onebox
