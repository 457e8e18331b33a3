#!/bin/bash
# round-2 multi-GPU call: usage  bash tools/gpu_r2_multi.sh N   (under gpurun --gpus N)
# bench.py weak + strong scaling, C5 (tools/pdf_bench.py) with the PDF hash that must equal the 1-GPU one
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 1200 $TR --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2m_bench_weak_${N}gpu.json 2> gpurun_out/r2m_bench_weak_${N}gpu.err
echo "weak rc=$?" >> gpurun_out/r2m_bench_weak_${N}gpu.err
timeout 1200 $TR --master-port 29518 bench.py --gpus $N --steps 5 --warmup 3 --no-alt --scaling strong --no-cpu-baseline --no-ref-gpu > gpurun_out/r2m_bench_strong_${N}gpu.json 2> gpurun_out/r2m_bench_strong_${N}gpu.err
echo "strong rc=$?" >> gpurun_out/r2m_bench_strong_${N}gpu.err
timeout 1200 $TR --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 --no-alt --geometry instanced --no-cpu-baseline --no-ref-gpu > gpurun_out/r2m_bench_weak_inst_${N}gpu.json 2> gpurun_out/r2m_bench_weak_inst_${N}gpu.err
echo "weak inst rc=$?" >> gpurun_out/r2m_bench_weak_inst_${N}gpu.err
EV=${2:-1000}
timeout 1200 $TR --master-port 29520 tools/pdf_bench.py --events $EV --instanced --out gpurun_out/r2m_c5_${EV}ev_inst_${N}gpu.json > gpurun_out/r2m_c5_inst_${N}gpu.log 2>&1
timeout 1200 $TR --master-port 29521 tools/pdf_bench.py --events $EV --out gpurun_out/r2m_c5_${EV}ev_flat_${N}gpu.json > gpurun_out/r2m_c5_flat_${N}gpu.log 2>&1
tail -n 3 gpurun_out/r2m_*_${N}gpu.err gpurun_out/r2m_c5_*_${N}gpu.log
