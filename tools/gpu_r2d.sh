#!/bin/bash
# round-2 GPU call D: config C4 on the 907-module detector (flat / two-level / out-of-line RNG draw), its ncu
# capture, config C5 on one GPU, small-stack variants on the two-level tree
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python tools/c4_bench.py --events 20 --out gpurun_out/r2d_c4_flat.json > gpurun_out/r2d_c4_flat.log 2>&1
timeout 900 python tools/c4_bench.py --events 20 --instanced --out gpurun_out/r2d_c4_inst.json > gpurun_out/r2d_c4_inst.log 2>&1
CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_rngool.so timeout 900 python tools/c4_bench.py --events 20 --out gpurun_out/r2d_c4_flat_rngool.json > gpurun_out/r2d_c4_flat_rngool.log 2>&1
CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_rngool.so timeout 600 python tools/prop_sweep.py '[{}]' 206 10000000 5 > gpurun_out/r2d_sweep_rngool.log 2>&1
for v in sstack8 sstack12; do
  CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200_$v.so timeout 600 python tools/prop_sweep.py '[{"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2d_sweep_inst_$v.log 2>&1
done
timeout 900 python tools/pdf_bench.py --events 200 --out gpurun_out/r2d_c5_1gpu_200ev_flat.json > gpurun_out/r2d_c5_flat.log 2>&1
timeout 900 python tools/pdf_bench.py --events 200 --instanced --out gpurun_out/r2d_c5_1gpu_200ev_inst.json > gpurun_out/r2d_c5_inst.log 2>&1
timeout 300 python tools/c4_bench.py --events 4 --oracle-sample 20000 > gpurun_out/r2d_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wide_propagate -s 6 -c 1 -f -o gpurun_out/r2d_c4_wide_prop \
    python tools/c4_bench.py --events 4 --oracle-sample 20000 > gpurun_out/r2d_ncu.log 2>&1
tail -n 3 gpurun_out/r2d_*.log
