#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_reference_gpu.py tests/test_gpu_instanced.py tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/r2stat_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2stat_tests.log
tail -n 6 gpurun_out/r2stat_tests.log
