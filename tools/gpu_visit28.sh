#!/bin/bash
# Re-rank kernel change: parity tests of the re-rank / evaluation path, then the short bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rerank.py tests/test_gpu_fullsize.py tests/test_gpu_dist.py -q -m gpu --timeout=300 -x > gpurun_out/t28.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/t28.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-assoc --no-cpu --no-retrieval > gpurun_out/bench28.json 2> gpurun_out/bench28.err; echo "bench rc=$?"; tail -3 gpurun_out/bench28.err
python -c "
import json;d=json.load(open('gpurun_out/bench28.json'));print(d['ms_per_step'],d['ms_per_step_stream_launches'],d['e2e']['value'],d['result']);print(d['kernel_share_of_step'], d['ms_per_step_with_kernel_events'])"
