#!/usr/bin/env bash
# Round-2 call 1 (gpurun --gpus 2): validation debt of round 1 -- full GPU suite, NVLS, topology, CLI job, 2-GPU bench.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
echo "== full GPU suite at $N GPUs"
timeout 900 python -u -m pytest tests -m gpu -q --timeout=200 -p no:cacheprovider > gpurun_out/pytest_gpu_all_${N}gpu.log 2>&1; echo "rc=$?"; tail -n 12 gpurun_out/pytest_gpu_all_${N}gpu.log | cut -c1-300
bash scripts/gpu_round2_first.sh 2>&1 | grep -v "multi-GPU suite" 
echo "== bench ours $N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 200 --warmup 20 2>&1 | grep -v Warning | tail -n 3 | tee gpurun_out/bench_ours_${N}_r2a.json | cut -c1-1500
