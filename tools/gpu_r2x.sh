#!/bin/bash
# round-2 GPU call X (1 GPU): guard moved to the ray-load sites — timing on both geometries, then the full suite and smoke()
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2x_c3.log 2>&1
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2x_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2x_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2x_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r2x_smoke.log
cut -c1-200 gpurun_out/r2x_c3.log; tail -n 3 gpurun_out/r2x_tests.log gpurun_out/r2x_smoke.log
