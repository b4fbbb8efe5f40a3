#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/frame_once.py 3 > gpurun_out/frame_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'crop_cv_linear_kernel|assoc_frame_kernel|det_pack_kernel' -s 3 -c 3 -o gpurun_out/r01_frame_full python tools/frame_once.py 2 > gpurun_out/ncu_frame.log 2>&1
echo "ncu frame rc=$?"; tail -3 gpurun_out/ncu_frame.log
