#!/bin/bash
# Run every GPU test module in its own process with a timeout (a hung kernel must not eat the box).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
rc=0
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary.txt
for t in tests/test_gpu_dist.py tests/test_gpu_crop.py tests/test_gpu_rerank.py tests/test_gpu_assoc.py tests/test_gpu_shard.py; do
  n=$(basename $t .py)
  timeout 600 python -m pytest $t -q -m gpu --timeout=300 > gpurun_out/$n.log 2>&1
  r=$?; echo "$n rc=$r" | tee -a gpurun_out/summary.txt
  [ $r -ne 0 ] && rc=1
  tail -5 gpurun_out/$n.log
done
tail -30 gpurun_out/smoke.log
exit $rc
