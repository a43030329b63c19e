#!/usr/bin/env bash
# 8 (or N) GPUs: the whole multi-GPU suite, soak, all-reduce sweep with NVLS, bench ours.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
echo "== multi-GPU suite at $N GPUs"
PDM_TEST_NVLS=1 timeout 700 python -u -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=200 -p no:cacheprovider > gpurun_out/pytest_multi_${N}gpu.log 2>&1; echo "rc=$?"; tail -n 8 gpurun_out/pytest_multi_${N}gpu.log | cut -c1-300
echo "== soak (skew 20 us, 20000 steps)"
PDM_SKEW_NS=20000 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566 benchmarks/soak.py --steps 20000 --out gpurun_out/soak_${N}.json 2>&1 | grep -v Warning | tail -n 2 | cut -c1-900
echo "== all-reduce sweep"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 benchmarks/allreduce_sweep.py --iters 50 --nvls --out gpurun_out/allreduce_nvls_${N}.json 2>&1 | grep -v Warning | tail -n 10 | cut -c1-420
echo "== bench ours $N"
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 200 --warmup 20 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_${N}_r2.json | cut -c1-4500
