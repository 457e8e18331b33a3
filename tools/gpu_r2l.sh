#!/bin/bash
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nproc > gpurun_out/r2l_probe.log; lscpu | grep -E "Model name|Socket|NUMA|Thread" >> gpurun_out/r2l_probe.log
timeout 600 python tools/pageable_probe.py >> gpurun_out/r2l_probe.log 2>&1
cat gpurun_out/r2l_probe.log
