#!/bin/bash
# Triangle-mode visit: its own tests, same-box A/B (REIDB200_TRI=0 = full matrix), then every GPU test, smoke, the bench
# line, the ncu launch list of the bench command and one --set full capture of a fused step.  If the mode's tests fail the
# rest of the visit runs with REIDB200_TRI=0 (the configuration that would ship then).
tag=${1:-tri}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu_$tag.txt 2>&1
timeout 200 python tools/tri_debug2.py > gpurun_out/tri_debug2_$tag.log 2>&1; grep -E "==|matrix:|cred differs|v1cols" gpurun_out/tri_debug2_$tag.log
timeout 420 python -m pytest tests/test_gpu_zy_triangle.py -q --timeout=300 > gpurun_out/t_tri_$tag.log 2>&1; trc=$?
echo "triangle tests rc=$trc" | tee gpurun_out/summary_$tag.txt; tail -15 gpurun_out/t_tri_$tag.log | cut -c1-300
timeout 300 bash tools/ab_kernels.sh REIDB200_TRI=0 2>&1 | tail -8
cp gpurun_out/ab_kernels.log gpurun_out/ab_kernels_$tag.log
if [ $trc -ne 0 ]; then export REIDB200_TRI=0; echo "REST OF THE VISIT WITH REIDB200_TRI=0" | tee -a gpurun_out/summary_$tag.txt; fi
timeout 900 python -m pytest tests -q -m gpu --timeout=600 --deselect tests/test_gpu_zy_triangle.py > gpurun_out/t_$tag.log 2>&1; echo "tests rc=$?" | tee -a gpurun_out/summary_$tag.txt
grep -E 'passed|failed|FAILED|ERROR' gpurun_out/t_$tag.log | tail -20 | cut -c1-300
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke_$tag.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/summary_$tag.txt
tail -2 gpurun_out/smoke_$tag.log | cut -c1-300
( time timeout 420 python bench.py > gpurun_out/bench_$tag.json 2> gpurun_out/bench_$tag.err ) 2>&1 | tail -3; echo "bench done" | tee -a gpurun_out/summary_$tag.txt
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_$tag.json").read().strip().splitlines()[-1])
    keep={k:d[k] for k in ("value","ms_per_step","e2e","roofline","kernel_share_of_step","survey_recipe","result","retrieval_ms_per_search") if k in d}
    print(json.dumps(keep)[:2500])
except Exception as e:
    print("no bench line", e)
PY
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench_$tag.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-retrieval --no-assoc --no-cpu-rows --no-sharded-rerank > gpurun_out/ncu_launch_$tag.log 2>&1; echo "launch list rc=$?" | tee -a gpurun_out/summary_$tag.txt
timeout 120 python tools/step_once_fused.py 2 > gpurun_out/plain_step_$tag.log 2>&1 && \
timeout 400 ncu --set full --clock-control none -s 15 -c 16 -f -o gpurun_out/step_full_$tag python tools/step_once_fused.py 2 > gpurun_out/ncu_step_$tag.log 2>&1; echo "ncu step rc=$?" | tee -a gpurun_out/summary_$tag.txt
ls -la gpurun_out | tail -12
