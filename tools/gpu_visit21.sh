#!/bin/bash
mkdir -p gpurun_out
for e in 0 1; do
if [ $e = 1 ]; then export REIDB200_BLOCKED_EXP=1; fi
timeout 300 python bench.py --steps 10 --warmup 3 --no-assoc --no-cpu --no-retrieval > gpurun_out/bench21_$e.json 2> gpurun_out/bench21_$e.err
python -c "
import json;d=json.load(open('gpurun_out/bench21_$e.json'));print($e, d['ms_per_step'], d['kernel_share_of_step']['gemm_split_tc_kernel']*d['ms_per_step'])"
done
