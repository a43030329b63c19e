#!/usr/bin/env bash
# Multi-GPU confirmation of the pulled LL exchange + flag dependency defaults (gpurun --gpus N).
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
run() { env "$@" STEPS=2000 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29681 benchmarks/step_sweep_multi.py 2>&1 | grep "us/step"; }
{
run PDM_NET_X=defaults
run PDM_NET_DEP=grid PDM_NET_LL=push
} | tee gpurun_out/step_sweep_multi_ll_$N.log
echo "== multi-gpu tests"
timeout 500 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider -k "ll_allreduce or native_cnn or full_job" > gpurun_out/pytest_multi_$N.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_multi_$N.log | cut -c1-300
echo "== bench ours N=$N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 1000 --warmup 50 > gpurun_out/bench_ours_$N.json 2> gpurun_out/bench_ours_$N.err; echo "rc=$?"; tail -1 gpurun_out/bench_ours_$N.json | cut -c1-900
