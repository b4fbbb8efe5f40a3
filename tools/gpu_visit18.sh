#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/step_once.py 2 > gpurun_out/step_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gemm_split -s 1 -c 1 -f -o gpurun_out/r01_gemm_sym_b python tools/step_once.py 2 > gpurun_out/ncu_gemm_b.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_gemm_b.log
