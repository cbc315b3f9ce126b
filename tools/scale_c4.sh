#!/bin/bash
# BASELINE config 4 (1024 clips, strong scaling) at N = 1, 2, 4, 8 ranks of one box + the default C2 line at N = 8.
# bash tools/scale_c4.sh TAG     (under gpurun --gpus 8)
TAG=${1:-c4}
O=gpurun_out; mkdir -p $O
for n in ${SRN_SCALE_NS:-1 2 4 8}; do
  if [ $n = 1 ]; then
    timeout 600 python bench.py --config C4 --core --steps 3 --warmup 3 > $O/${TAG}_c4_n$n.json 2> $O/${TAG}_c4_n$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n \
      bench.py --config C4 --core --gpus $n --steps 3 --warmup 3 > $O/${TAG}_c4_n$n.json 2> $O/${TAG}_c4_n$n.err
  fi
  echo "C4 n=$n rc=$?"; python - <<PY
import json
try:
    d = json.loads(open("$O/${TAG}_c4_n$n.json").read().strip().splitlines()[-1])
    print("  value %.0f f/s  e2e %.0f f/s  ms/step %.1f  checksum %.6f  clocks %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["config"]["checksum_all_mels"], d["clocks"]))
except Exception as e:
    print("  no line:", e)
PY
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 \
  bench.py --gpus 8 --core --steps 10 --warmup 3 > $O/${TAG}_c2_n8.json 2> $O/${TAG}_c2_n8.err; echo "C2 n=8 rc=$?"; tail -c 600 $O/${TAG}_c2_n8.json
