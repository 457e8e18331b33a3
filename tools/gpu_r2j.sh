#!/bin/bash
# round-2 GPU call J (1 GPU): predicated fast push (WIDE_FAST_PUSH) A/B on C3 and C4, flat and two-level
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for v in "" _nofp _tp _tps; do
  export CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200$v.so
  timeout 600 python tools/prop_sweep.py '[{}, {"_instanced": 1}]' 206 10000000 5 > gpurun_out/r2j_c3$v.log 2>&1
  timeout 600 python tools/c4_bench.py --events 8 --oracle-sample 20000 --out gpurun_out/r2j_c4_flat$v.json > gpurun_out/r2j_c4_flat$v.log 2>&1
  timeout 600 python tools/c4_bench.py --events 8 --oracle-sample 20000 --instanced --out gpurun_out/r2j_c4_inst$v.json > gpurun_out/r2j_c4_inst$v.log 2>&1
done
unset CHROMA_B200_LIB
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_intersect.py tests/test_gpu_propagate.py tests/test_gpu_instanced.py tests/test_gpu_render.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2j_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2j_tests.log
tail -n 2 gpurun_out/r2j_c3*.log; tail -n 3 gpurun_out/r2j_tests.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2j_c4_*.json')):
    d=json.load(open(f)); print(f, d['propagate_launch_ms'], d['photons_per_s_device_sources'], d['oracle_flags_identical'])
PY
