#!/bin/bash
# round-2 GPU call B: full GPU suite on the product library, A/B-kernel tests on the ab variant,
# kernel variants A/B (two-pass group step, leaf spread), first numbers of the two-level geometry
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2b_tests.log
CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_ab.so timeout 600 python -m pytest tests/test_gpu_wide.py -q > gpurun_out/r2b_tests_ab.log 2>&1
echo "ab rc=$?" >> gpurun_out/r2b_tests_ab.log
timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}, {"_instanced": 1, "CHROMA_REFILL": 12}, {"_instanced": 1, "CHROMA_REFILL": 26}, {"_instanced": 1, "CHROMA_WIDE_LEAF_THRESH": 4}, {"_instanced": 1, "CHROMA_WIDE_LEAF_THRESH": 12}]' 206 10000000 5 > gpurun_out/r2b_sweep.log 2>&1
for v in twopass twopass2 spread spread2p; do
  CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_$v.so timeout 600 python tools/prop_sweep.py '[{}]' 206 10000000 5 > gpurun_out/r2b_sweep_$v.log 2>&1
done
CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_inst3.so timeout 600 python tools/prop_sweep.py '[{"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2b_sweep_inst3.log 2>&1
tail -n 4 gpurun_out/r2b_*.log
