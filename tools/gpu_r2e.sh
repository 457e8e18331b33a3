#!/bin/bash
# round-2 GPU call E: full suite, flat / two-level sweeps with the batched enters + leaf spread, C4 with the
# out-of-line-RNG kernel build picked by geometry, a full bench.py line + the reference arm
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2e_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2e_tests.log
timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}, {"_instanced": 1, "CHROMA_WIDE_LEAF_THRESH": 6}, {"_instanced": 1, "CHROMA_WIDE_LEAF_THRESH": 12}, {"_instanced": 1, "CHROMA_REFILL": 16}, {"_instanced": 1, "CHROMA_REFILL": 24}]' 206 10000000 5 > gpurun_out/r2e_sweep.log 2>&1
timeout 900 python tools/c4_bench.py --events 20 --out gpurun_out/r2e_c4_flat.json > gpurun_out/r2e_c4_flat.log 2>&1
timeout 900 python tools/c4_bench.py --events 20 --instanced --out gpurun_out/r2e_c4_inst.json > gpurun_out/r2e_c4_inst.log 2>&1
timeout 1200 python bench.py > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err
echo "bench rc=$?" >> gpurun_out/r2e_bench.err
timeout 1200 python bench.py --impl reference > gpurun_out/r2e_bench_reference.json 2> gpurun_out/r2e_bench_reference.err
echo "ref rc=$?" >> gpurun_out/r2e_bench_reference.err
tail -n 4 gpurun_out/r2e_tests.log gpurun_out/r2e_sweep.log gpurun_out/r2e_bench.err gpurun_out/r2e_bench_reference.err
