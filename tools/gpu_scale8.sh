#!/bin/bash
# N-GPU host-ceiling visit: concurrent pinned H2D bandwidth of N ranks (unbound / NUMA-bound), then the bench line at N ranks
# with and without the binding (headline + e2e only).   usage: bash tools/gpu_scale8.sh N
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
lscpu | grep -E "Model name|Socket|NUMA|^CPU\(s\)" > gpurun_out/lscpu_$N.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29521 tools/h2d_bw_ranks.py > gpurun_out/h2d_ranks_$N.json 2> gpurun_out/h2d_ranks_$N.err
echo "h2d rc=$?"; grep '^{' gpurun_out/h2d_ranks_$N.json
MODES=("")
grep -q '"bound": true' gpurun_out/h2d_ranks_$N.json && MODES=("" "--no-numa-bind")   # the second arm only where binding is possible
for mode in "${MODES[@]}"; do
  tag=n${N}$( [ -n "$mode" ] && echo _unbound )
  timeout 600 $TR --master-port 29522 bench.py --gpus $N --steps 20 --warmup 3 --no-assoc --no-cpu --no-retrieval --no-cpu-rows --no-sharded-rerank $mode > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err
  echo "bench $tag rc=$?"
  grep '^{' gpurun_out/bench_$tag.json | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('value %.4g ms %.4f e2e %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value']), d['e2e'].get('host_numa'))"
done
