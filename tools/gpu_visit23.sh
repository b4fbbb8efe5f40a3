#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_posttricks.py -q -m gpu --timeout=120 -x > gpurun_out/t23.log 2>&1; echo "tests rc=$?"; tail -15 gpurun_out/t23.log
timeout 200 python tools/dbscan_once.py 3 2>&1 | tail -3
