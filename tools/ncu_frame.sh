#!/bin/bash
# ncu --set full of the frame kernels (MOT20 shape).  usage: ncu_frame.sh <tag> [env assignments...]
tag=$1; shift
mkdir -p gpurun_out
env "$@" timeout 600 python tools/frame_once.py 3 > gpurun_out/frame_plain_$tag.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/frame_plain_$tag.log; exit 1; }
env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k regex:'assoc_frame|crop_cv|det_pack' -s 3 -c 6 -f -o gpurun_out/frame_$tag python tools/frame_once.py 3 > gpurun_out/ncu_frame_$tag.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_frame_$tag.log
python tools/ncu_summary.py gpurun_out/frame_$tag.ncu-rep > gpurun_out/frame_${tag}_summary.csv 2>/dev/null; cat gpurun_out/frame_${tag}_summary.csv | cut -c1-400
