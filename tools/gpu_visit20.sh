#!/bin/bash
mkdir -p gpurun_out
for k in "$@"; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 1 -c 1 -f -o gpurun_out/r01_$k python tools/step_once.py 2 > gpurun_out/ncu_$k.log 2>&1
echo "$k rc=$?"
done
