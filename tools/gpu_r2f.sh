#!/bin/bash
# round-2 GPU call F: A/B of (two-pass, leaf spread, batched enters) on C3 AND C4, flat and two-level
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for v in "" _v00 _v01 _v10 _ibe; do
  export CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200$v.so
  timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2f_c3$v.log 2>&1
  timeout 600 python tools/c4_bench.py --events 8 --oracle-sample 20000 --out gpurun_out/r2f_c4_flat$v.json > gpurun_out/r2f_c4_flat$v.log 2>&1
  timeout 600 python tools/c4_bench.py --events 8 --oracle-sample 20000 --instanced --out gpurun_out/r2f_c4_inst$v.json > gpurun_out/r2f_c4_inst$v.log 2>&1
done
unset CHROMA_B200_LIB
tail -n 2 gpurun_out/r2f_c3*.log
