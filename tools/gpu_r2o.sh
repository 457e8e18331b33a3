#!/bin/bash
# round-2 GPU call O (1 GPU): ncu --set full of wide_intersect_kernel at 1e7 rays (the command first runs without ncu)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
CMD="python tools/rays_probe.py 10000000"
timeout 600 $CMD > gpurun_out/r2o_plain.log 2>&1
if [ $? -eq 0 ]; then
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:wide_intersect -s 2 -c 1 -f -o gpurun_out/r2o_wide_isect $CMD > gpurun_out/r2o_ncu.log 2>&1
  echo "ncu rc=$?" >> gpurun_out/r2o_ncu.log
fi
tail -n 3 gpurun_out/r2o_plain.log gpurun_out/r2o_ncu.log
