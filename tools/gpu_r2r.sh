#!/bin/bash
# round-2 GPU call R (1 GPU): world-frame slab constants saved across placement visits (INST_SAVE_FRAME) — same-box A/B
# on C3 and C4 (two-level geometry), then the two-level tests
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
for v in "" _nosf; do
  CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200$v.so timeout 400 python tools/prop_sweep.py tools/inst_only.json 206 10000000 7 > gpurun_out/r2r_c3$v.log 2>&1
  CHROMA_B200_LIB=$PWD/chroma_b200/libchroma_b200$v.so timeout 400 python tools/c4_bench.py --events 8 --oracle-sample 20000 --instanced --out gpurun_out/r2r_c4_inst$v.json > gpurun_out/r2r_c4_inst$v.log 2>&1
done
timeout 600 python -m pytest tests/test_gpu_instanced.py tests/test_gpu_wide.py -m gpu -q > gpurun_out/r2r_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2r_tests.log
cut -c1-170 gpurun_out/r2r_c3*.log; tail -n 3 gpurun_out/r2r_tests.log
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/r2r_c4_*.json')):
    d=json.load(open(f)); print(f, d['propagate_launch_ms'], d['photons_per_s_device_sources'], d['oracle_flags_identical'])
PY
