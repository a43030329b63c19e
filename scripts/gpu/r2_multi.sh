#!/usr/bin/env bash
# N GPUs: multi-GPU suite (persistent tensor-core kernel included), bench at N.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
timeout 200 python -u -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout=100 -p no:cacheprovider -k "persistent_tc" 2>&1 | tail -n 15 | cut -c1-300
echo "== full multi suite"
PDM_TEST_NVLS=1 timeout 600 python -u -m pytest tests/test_gpu_multi.py -m gpu -q --timeout=150 -p no:cacheprovider > gpurun_out/pytest_multi_${N}gpu.log 2>&1; echo "rc=$?"; tail -n 6 gpurun_out/pytest_multi_${N}gpu.log | cut -c1-300
echo "== bench ours $N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 200 --warmup 20 2>&1 | grep -v Warning | tail -n 2 | tee gpurun_out/bench_ours_${N}_tc.json | cut -c1-3500
