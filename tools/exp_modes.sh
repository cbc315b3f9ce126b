for m in 0 1 2 3 4 7; do
  SRN_DBG_MODE=$m python tools/profile_ops.py 16 1000 gpurun_out/exp_mode$m.txt > /dev/null 2>&1
  echo "mode $m: $(head -1 gpurun_out/exp_mode$m.txt)"
done
