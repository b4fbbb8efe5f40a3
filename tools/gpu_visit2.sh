#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard.py -q -m gpu --timeout=300 > gpurun_out/test_gpu_shard.log 2>&1; echo "test_gpu_shard rc=$?" | tee -a gpurun_out/summary2.txt
tail -15 gpurun_out/test_gpu_shard.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench2.json 2> gpurun_out/bench2.err; echo "bench rc=$?" | tee -a gpurun_out/summary2.txt
tail -5 gpurun_out/bench2.err; cat gpurun_out/bench2.json
timeout 300 python tools/step_once.py 2 > gpurun_out/step_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -s 13 -c 13 -o gpurun_out/r01_step_full python tools/step_once.py 2 > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?" | tee -a gpurun_out/summary2.txt
tail -3 gpurun_out/ncu_full.log
ls -la gpurun_out
