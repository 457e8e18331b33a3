#!/bin/bash
# round-2 GPU call V (N GPUs): bench.py with the DRIVER's flags under torchrun (rank-0-only legs must stay out of collectives)
N=${1:-2}
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29531 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2v_bench_default_${N}gpu.json 2> gpurun_out/r2v_bench_default_${N}gpu.err
echo "default rc=$?" >> gpurun_out/r2v_bench_default_${N}gpu.err
timeout 300 $TR --master-port 29532 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/r2v_bench_reference_${N}gpu.json 2> gpurun_out/r2v_bench_reference_${N}gpu.err
echo "reference rc=$?" >> gpurun_out/r2v_bench_reference_${N}gpu.err
tail -n 3 gpurun_out/r2v_bench_default_${N}gpu.err gpurun_out/r2v_bench_reference_${N}gpu.err; cut -c1-300 gpurun_out/r2v_bench_default_${N}gpu.json
