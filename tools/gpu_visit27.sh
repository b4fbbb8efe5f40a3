#!/bin/bash
# Query-sharded evaluate_all: GPU test at world size 1, torchrun check + bench at 2 ranks.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_shard.py -q -m gpu --timeout=300 -x > gpurun_out/t27.log 2>&1; echo "tests rc=$?"; tail -5 gpurun_out/t27.log
bash tools/gpu_multi.sh 2
cat gpurun_out/shard_check_2.log | tail -12
python -c "
import json
for l in open('gpurun_out/bench_n2.json'):
    if l.startswith('{'): print(json.dumps(json.loads(l)['eval_c1']))"
timeout 600 python bench.py --no-assoc --no-cpu --no-retrieval > gpurun_out/bench27.json 2> gpurun_out/bench27.err; echo "bench rc=$?"; tail -3 gpurun_out/bench27.err
python -c "
import json;d=json.load(open('gpurun_out/bench27.json'));print(d['ms_per_step'], json.dumps(d['eval_c1']))"
