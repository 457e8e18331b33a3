#!/bin/bash
# round-2 GPU call N (1 GPU): borrowed pageable uploads (tests + bench variants + C4 host bombs), C5 with 1000 events on one GPU
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hygiene.py tests/test_gpu_daq_pdf.py tests/test_gpu_propagate.py tests/test_gpu_dist.py tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r2n_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2n_tests.log
timeout 1200 python bench.py --no-cpu-baseline --no-ref-gpu > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err
echo "bench rc=$?" >> gpurun_out/r2n_bench.err
timeout 900 python tools/c4_bench.py --events 20 --out gpurun_out/r2n_c4_flat.json > gpurun_out/r2n_c4_flat.log 2>&1
timeout 900 python tools/pdf_bench.py --events 1000 --out gpurun_out/r2n_c5_1000ev_flat_1gpu.json > gpurun_out/r2n_c5_flat.log 2>&1
timeout 900 python tools/pdf_bench.py --events 1000 --instanced --out gpurun_out/r2n_c5_1000ev_inst_1gpu.json > gpurun_out/r2n_c5_inst.log 2>&1
tail -n 5 gpurun_out/r2n_tests.log gpurun_out/r2n_bench.err
python - <<'PY'
import json
b=json.loads(open('gpurun_out/r2n_bench.json').read().strip().split('\n')[-1])
print(b['value'], b['e2e']['value'], json.dumps(b['e2e_variants']))
d=json.load(open('gpurun_out/r2n_c4_flat.json')); print('c4', d['photons_per_s_device_sources'], d['photons_per_s_host_bombs'])
for f in ('flat','inst'):
    d=json.load(open('gpurun_out/r2n_c5_1000ev_%s_1gpu.json'%f)); print(f, d['value'], d['pdf_sha256'][:16], d['hitcount_equals_pdf_sum'])
PY
