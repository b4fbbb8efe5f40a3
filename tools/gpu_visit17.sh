#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -q -m gpu --timeout=300 -x > gpurun_out/t17.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t17.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-assoc --no-cpu --retrieval-gallery 1000000 > gpurun_out/bench17.json 2> gpurun_out/bench17.err; echo "bench rc=$?"
tail -5 gpurun_out/bench17.err; python -c "
import json;d=json.load(open('gpurun_out/bench17.json'));print(d['value'],d['ms_per_step'],d['roofline']['frac'],d['kernel_share_of_step']);r=d['retrieval'];print(r['ms_per_search'],r['gemm_frac_of_peak_div3'],r['kernel_ms'])"
