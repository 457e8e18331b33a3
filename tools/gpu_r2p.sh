#!/bin/bash
# round-2 GPU call P (2 GPUs): chroma_daq_pack / chroma_daq_merge_gathered — the GPU test, then the strong-scaling bench
# (reduce_channels over NCCL inside the timed region) and the channel merge checked against one GPU
cd "$GRAFT_REPO_ROOT" 2>/dev/null || true
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_daq_pdf.py tests/test_gpu_dist.py -m gpu -q > gpurun_out/r2p_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r2p_tests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 420 $TR --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 --no-alt --scaling strong --no-cpu-baseline --no-ref-gpu > gpurun_out/r2p_bench_strong_2gpu.json 2> gpurun_out/r2p_bench_strong_2gpu.err
echo "strong rc=$?" >> gpurun_out/r2p_bench_strong_2gpu.err
timeout 300 $TR --master-port 29542 tools/strong_check.py > gpurun_out/r2p_strong_check.log 2>&1
echo "check rc=$?" >> gpurun_out/r2p_strong_check.log
tail -n 3 gpurun_out/r2p_tests.log gpurun_out/r2p_bench_strong_2gpu.err gpurun_out/r2p_strong_check.log; cut -c1-200 gpurun_out/r2p_bench_strong_2gpu.json
