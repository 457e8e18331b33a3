#!/bin/bash
# round-2 GPU call C: full GPU suite (kernel defaults: two-pass group step + leaf spread), opened placements
# of the two-level tree, ncu capture of the two-level propagate kernel, a first run of the new bench.py
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2c_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2c_tests.log
timeout 600 python tools/prop_sweep.py '[{"_instanced": 1}, {"_instanced": 1, "CHROMA_INST_OPEN": 0}, {"_instanced": 1, "CHROMA_INST_OPEN": 2}, {"_instanced": 1, "CHROMA_INST_OPEN": 5}]' 206 10000000 5 > gpurun_out/r2c_sweep_inst.log 2>&1
timeout 600 python tools/prop_sweep.py '[{}]' 206 10000000 5 > gpurun_out/r2c_sweep_flat.log 2>&1
timeout 900 python bench.py --steps 3 --warmup 3 --cpu-seconds 5 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err
echo "bench rc=$?" >> gpurun_out/r2c_bench.err
timeout 300 python tools/prop_sweep.py '[{"_instanced": 1}]' 206 10000000 1 > gpurun_out/r2c_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:inst_propagate -s 1 -c 1 -f -o gpurun_out/r2c_inst_prop \
    python tools/prop_sweep.py '[{"_instanced": 1}]' 206 10000000 1 > gpurun_out/r2c_ncu.log 2>&1
tail -n 5 gpurun_out/r2c_tests.log gpurun_out/r2c_sweep_*.log gpurun_out/r2c_bench.err
