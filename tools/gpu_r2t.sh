#!/bin/bash
# round-2 GPU call T (1 GPU): same box A/B — empty-index filters (default) against the per-child empty-record test
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for k in 1 2; do
  for v in "" _ec1; do
    CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200$v.so timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}]' 206 10000000 7 > gpurun_out/r2t_c3${v}_$k.log 2>&1
  done
done
for f in gpurun_out/r2t_c3*.log; do echo $f; cut -c1-120 $f; done
