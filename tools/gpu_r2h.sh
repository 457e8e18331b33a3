#!/bin/bash
# round-2 GPU call H (1 GPU): area-cost splits in the wide-tree builder (CHROMA_WIDE_SAH = largest part split that way)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1200 python tools/prop_sweep.py '[{}, {"CHROMA_WIDE_SAH": 64}, {"CHROMA_WIDE_SAH": 4096}, {"CHROMA_WIDE_SAH": 65536}, {"_instanced": 1}, {"_instanced": 1, "CHROMA_WIDE_SAH": 4096}]' 206 10000000 5 > gpurun_out/r2h_sweep.log 2>&1
echo "sweep rc=$?" >> gpurun_out/r2h_sweep.log
CHROMA_WIDE_SAH=4096 timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_intersect.py tests/test_gpu_propagate.py tests/test_gpu_instanced.py tests/test_gpu_render.py -m gpu -q > gpurun_out/r2h_tests_sah.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2h_tests_sah.log
CHROMA_WIDE_SAH=4096 timeout 600 python tools/c4_bench.py --events 8 --oracle-sample 20000 --out gpurun_out/r2h_c4_flat_sah.json > gpurun_out/r2h_c4_flat_sah.log 2>&1
tail -n 8 gpurun_out/r2h_sweep.log; tail -n 3 gpurun_out/r2h_tests_sah.log
