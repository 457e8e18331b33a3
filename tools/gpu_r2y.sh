#!/bin/bash
# round-2 GPU call Y (1 GPU): call Z again after the unsearchable-ray guard (no reference arm, no leaf sweep)
# (engine, reference arm, two-level as headline), C4 both geometries, leaf-size check, launch list of the bench command
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2y_noalloc_alone.log 2>&1
echo "noalloc alone rc=$?" >> gpurun_out/r2y_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r2y_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2y_tests.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2y_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/r2y_smoke.log
timeout 1200 python bench.py > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err
echo "bench rc=$?" >> gpurun_out/r2y_bench.err
timeout 1200 python bench.py --geometry instanced --no-alt --no-cpu-baseline --no-ref-gpu > gpurun_out/r2y_bench_two_level.json 2> gpurun_out/r2y_bench_two_level.err
timeout 900 python tools/c4_bench.py --events 20 --out gpurun_out/r2y_c4_flat.json > gpurun_out/r2y_c4_flat.log 2>&1
timeout 900 python tools/c4_bench.py --events 20 --instanced --out gpurun_out/r2y_c4_inst.json > gpurun_out/r2y_c4_inst.log 2>&1
B="python bench.py --steps 2 --warmup 3 --no-alt --no-cpu-baseline --no-ref-gpu"
timeout 900 $B > gpurun_out/r2y_bench_short.json 2> gpurun_out/r2y_bench_short.err
if [ $? -eq 0 ]; then
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2y_launches.csv $B > gpurun_out/r2y_ncu_list.log 2>&1
  echo "ncu rc=$?" >> gpurun_out/r2y_ncu_list.log
fi
tail -n 4 gpurun_out/r2y_noalloc_alone.log gpurun_out/r2y_tests.log gpurun_out/r2y_smoke.log gpurun_out/r2y_bench.err
