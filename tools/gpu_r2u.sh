#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for k in 1 2 3 4 5 6; do timeout 300 python -m pytest tests/test_reference_gpu.py -k "scintillator" -m gpu -q > gpurun_out/r2u_$k.log 2>&1; echo "rc=$?" >> gpurun_out/r2u_$k.log; done
grep -h "passed\|failed\|^E   " gpurun_out/r2u_*.log | cut -c1-160
