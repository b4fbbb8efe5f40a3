#!/bin/bash
# N-GPU visit: the multi-GPU tests, the NCCL paths of the sharded search / sharded re-ranked evaluation (parity) and the
# bench line at N ranks.   usage: bash tools/gpu_multi.sh N [bench args]
N=${1:-2}; shift
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpus_$N.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_shard.py tests/test_gpu_shims.py -q -m gpu --timeout=600 > gpurun_out/t_multi_$N.log 2>&1
echo "multi tests N=$N rc=$?" | tee -a gpurun_out/summary_multi.txt
grep -E 'passed|failed|FAILED|ERROR' gpurun_out/t_multi_$N.log | tail -10
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tests/nccl_shard_check.py > gpurun_out/shard_check_$N.log 2>&1
echo "shard_check N=$N rc=$?" | tee -a gpurun_out/summary_multi.txt
grep shard_check gpurun_out/shard_check_$N.log | tail -12
grep -E 'Error|error|Traceback' gpurun_out/shard_check_$N.log | head -5
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench N=$N rc=$?" | tee -a gpurun_out/summary_multi.txt
tail -3 gpurun_out/bench_n$N.err; grep '^{' gpurun_out/bench_n$N.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('value',d['value'],'ms',d['ms_per_step'],'e2e',d['e2e']['value']); print(json.dumps(d.get('rerank_sharded'),indent=1)); print(json.dumps(d.get('retrieval'))[:1500])"
