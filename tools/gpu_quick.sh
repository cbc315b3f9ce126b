#!/bin/bash
# Short GPU pass: parity tests, timing of C1/C2 solves, per-op table.  bash tools/gpu_quick.sh TAG
TAG=${1:-q}
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $O/${TAG}_pytest.log
timeout 300 python tests/diagnose.py --time > $O/${TAG}_diag_bf16.log 2>&1; echo "diag rc=$?"; grep "B=" $O/${TAG}_diag_bf16.log
timeout 300 python tools/profile_ops.py 16 1000 $O/${TAG}_ops_c2.txt > /dev/null 2>&1; echo "ops rc=$?"; head -1 $O/${TAG}_ops_c2.txt
