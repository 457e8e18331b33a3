#!/bin/bash
# round-2 GPU call Q (1 GPU): the final build's hygiene / full-size files three more times in fresh processes
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for k in 1 2 3; do timeout 600 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_hygiene.py tests/test_gpu_wide.py -m gpu -q > gpurun_out/r2q_hyg$k.log 2>&1; echo "rc=$?" >> gpurun_out/r2q_hyg$k.log; done
tail -n 3 gpurun_out/r2q_hyg*.log
