#!/bin/bash
# round-2 GPU call I (1 GPU): ncu --set full of the final flat propagate kernel at C3 (the command first runs without ncu)
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
CMD="python tools/prop_sweep.py tools/inst_only.json 206 10000000 1"
timeout 600 $CMD > gpurun_out/r2s_plain.log 2>&1
if [ $? -eq 0 ]; then
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:inst_propagate -s 1 -c 1 -f -o gpurun_out/r2s_inst_prop $CMD > gpurun_out/r2s_ncu.log 2>&1
  echo "ncu rc=$?" >> gpurun_out/r2s_ncu.log
fi
tail -n 3 gpurun_out/r2s_plain.log gpurun_out/r2s_ncu.log
