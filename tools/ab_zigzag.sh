#!/bin/bash
# A/B of the zig-zag tile order (SRN_ZIGZAG=0 / 1) on one box: C2 and C5 solves, alternating, plus the checksum of a
# solve in both orders (must be identical: the order of the tiles does not enter any value).
TAG=${1:-zz}; O=gpurun_out; mkdir -p $O
for rep in 1 2 3; do
  for z in 0 1; do
    SRN_ZIGZAG=$z timeout 200 python bench.py --core --steps 20 --warmup 5 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('C2 zigzag=$z', round(d['ms_per_step'],3), 'ms')" | tee -a $O/${TAG}_ab.txt
  done
done
for z in 0 1; do
  SRN_ZIGZAG=$z timeout 300 python bench.py --config C5 --core --steps 3 --warmup 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('C5 zigzag=$z', round(d['ms_per_step'],2), 'ms')" | tee -a $O/${TAG}_ab.txt
  SRN_ZIGZAG=$z timeout 300 python bench.py --config C3 --core --steps 3 --warmup 1 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('C3 zigzag=$z', round(d['ms_per_step'],2), 'ms')" | tee -a $O/${TAG}_ab.txt
  SRN_ZIGZAG=$z timeout 200 python tools/determinism_solve.py 3 2>&1 | tail -2 | tee -a $O/${TAG}_ab.txt
done
