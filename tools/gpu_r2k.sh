#!/bin/bash
# round-2 GPU call K (1 GPU): empty-record test dropped (A/B), knob sweep on the new kernels
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
S='[{}, {"_instanced": 1}, {"CHROMA_WIDE_LEAF_THRESH": 4}, {"CHROMA_WIDE_LEAF_THRESH": 6}, {"CHROMA_WIDE_LEAF_THRESH": 12}, {"CHROMA_REFILL": 16}, {"CHROMA_REFILL": 24}, {"CHROMA_REFILL": 28}, {"CHROMA_WIDE_ITERS": 8}, {"CHROMA_WIDE_ITERS": 32}, {"_instanced": 1, "CHROMA_WIDE_LEAF_THRESH": 5}, {"_instanced": 1, "CHROMA_WIDE_LEAF_THRESH": 12}, {"_instanced": 1, "CHROMA_REFILL": 16}, {"_instanced": 1, "CHROMA_REFILL": 24}]'
timeout 900 python tools/prop_sweep.py "$S" 206 10000000 5 > gpurun_out/r2k_sweep.log 2>&1
CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_ec.so timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2k_c3_ec.log 2>&1
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_intersect.py tests/test_gpu_propagate.py tests/test_gpu_instanced.py tests/test_gpu_render.py tests/test_gpu_fullsize.py -m gpu -q > gpurun_out/r2k_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2k_tests.log
cat gpurun_out/r2k_sweep.log | cut -c1-200; tail -n 2 gpurun_out/r2k_c3_ec.log | cut -c1-200; tail -n 3 gpurun_out/r2k_tests.log
