#!/bin/bash
# usage: gpu_ncu_fused.sh <kernel-regex> <out-name> [skip]   -- ncu --set full of one kernel of the fused hard-data step
mkdir -p gpurun_out
timeout 300 python tools/step_once_fused.py 2 > gpurun_out/plain_$2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s ${3:-1} -c 1 -f -o gpurun_out/$2 python tools/step_once_fused.py 2 > gpurun_out/ncu_$2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/plain_$2.log; tail -2 gpurun_out/ncu_$2.log
