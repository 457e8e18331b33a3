#!/bin/bash
# round-2 GPU call A: new parity tests + two-pass group step A/B
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_smi.txt
timeout 900 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_daq_pdf.py tests/test_reference_gpu.py -x -q > gpurun_out/r2a_tests_small.log 2>&1
echo "small rc=$?" >> gpurun_out/r2a_tests_small.log
timeout 1200 python -m pytest tests/test_gpu_fullsize.py -x -q > gpurun_out/r2a_tests_full.log 2>&1
echo "full rc=$?" >> gpurun_out/r2a_tests_full.log
for v in "" _twopass _twopass2; do
  CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200$v.so timeout 600 python tools/prop_sweep.py '[{}]' 206 10000000 5 > gpurun_out/r2a_sweep$v.log 2>&1
done
tail -3 gpurun_out/r2a_tests_small.log gpurun_out/r2a_tests_full.log gpurun_out/r2a_sweep*.log
