#!/bin/bash
# round-2 final GPU call (1 GPU): the committed tree once more — no-alloc test alone, full GPU suite, smoke(), a short bench line
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2final_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?" >> gpurun_out/r2final_noalloc_alone.log
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/r2final_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2final_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2final_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r2final_smoke.log
tail -n 3 gpurun_out/r2final_noalloc_alone.log gpurun_out/r2final_tests.log gpurun_out/r2final_smoke.log
