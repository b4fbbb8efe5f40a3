#!/bin/bash
# usage: gpu_ncu_kernel.sh <kernel-regex> <out-name> [skip]
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$1 -s ${3:-1} -c 1 -f -o gpurun_out/$2 python tools/step_once.py 2 > gpurun_out/ncu_$2.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_$2.log
