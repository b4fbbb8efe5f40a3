#!/bin/bash
# N-GPU visit: NCCL path of the sharded search (parity) + bench at N ranks.   usage: bash tools/gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_$N.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/nccl_shard_check.py > gpurun_out/shard_check_$N.log 2>&1
echo "shard_check N=$N rc=$?" | tee -a gpurun_out/summary_multi.txt
grep shard_check gpurun_out/shard_check_$N.log | tail -5
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$?" | tee -a gpurun_out/summary_multi.txt
tail -3 gpurun_out/bench_n$N.err; grep '^{' gpurun_out/bench_n$N.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value']); print(json.dumps(d.get('retrieval'),indent=1))"
