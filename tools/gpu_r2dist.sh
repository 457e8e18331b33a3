#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_dist.py -m gpu -q > gpurun_out/r2dist_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2dist_tests.log
tail -n 15 gpurun_out/r2dist_tests.log
