#!/bin/bash
# Round-2 GPU pass: parity tests, the bench line (all legs), per-op table, clock64 phase trace (trace variant).
# bash tools/gpu_r2.sh TAG [configs...]
TAG=${1:-r02}; shift
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $O/${TAG}_pytest.log
timeout 600 python bench.py > $O/${TAG}_bench_C2.json 2> $O/${TAG}_bench_C2.err; echo "bench C2 rc=$?"; tail -c 3000 $O/${TAG}_bench_C2.json; tail -3 $O/${TAG}_bench_C2.err
for c in "$@"; do
  timeout 900 python bench.py --config $c --steps 3 > $O/${TAG}_bench_$c.json 2> $O/${TAG}_bench_$c.err; echo "bench $c rc=$?"; tail -c 1500 $O/${TAG}_bench_$c.json; tail -3 $O/${TAG}_bench_$c.err
done
timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_c2.txt > /dev/null 2>&1; echo "ops rc=$?"; head -1 $O/${TAG}_ops_c2.txt
if [ -f serenade_b200/libserenade_b200_trace.so ]; then
  SRN_LIB_VARIANT=trace SRN_TRACE=1 timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_trace.txt > /dev/null 2> $O/${TAG}_trace.txt; echo "trace rc=$?"; grep -c trace $O/${TAG}_trace.txt
fi
