#!/bin/bash
mkdir -p gpurun_out
for dbg in 0 1 2 4; do
REIDB200_GEMM_DEBUG=$dbg timeout 300 python bench.py --steps 10 --warmup 3 --no-assoc --no-cpu --no-retrieval > gpurun_out/bench19_$dbg.json 2> gpurun_out/bench19_$dbg.err
python -c "
import json;d=json.load(open('gpurun_out/bench19_$dbg.json'));print($dbg, d['ms_per_step'], d['kernel_share_of_step']['gemm_split_tc_kernel']*d['ms_per_step'])"
done
