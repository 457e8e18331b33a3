#!/bin/bash
# round-2 GPU call W (1 GPU): after "a stalled upload is not read" + empty-index filters: timing, the hygiene file
# three times in fresh processes, the full suite, smoke()
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2w_c3.log 2>&1
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2w_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2w_tests.log
for k in 1 2 3; do timeout 600 python -m pytest tests/test_gpu_fullsize.py tests/test_gpu_hygiene.py -m gpu -q > gpurun_out/r2w_hyg$k.log 2>&1; echo "rc=$?" >> gpurun_out/r2w_hyg$k.log; done
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2w_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r2w_smoke.log
cut -c1-200 gpurun_out/r2w_c3.log; tail -n 3 gpurun_out/r2w_tests.log gpurun_out/r2w_hyg*.log gpurun_out/r2w_smoke.log
