#!/bin/bash
# One GPU box visit: smoke + GPU tests, bench (ours + reference arm), ncu launch list of the re-ranked step.
mkdir -p gpurun_out
bash tools/gpu_check.sh
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?" | tee -a gpurun_out/summary.txt
tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/bench_ref.json
timeout 300 python tools/step_once.py 3 > gpurun_out/step_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python tools/step_once.py 3 > gpurun_out/ncu_launch.log 2>&1
echo "ncu launches rc=$?" | tee -a gpurun_out/summary.txt
cat gpurun_out/summary.txt
