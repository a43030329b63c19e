#!/usr/bin/env bash
# Multi-GPU pass (gpurun --gpus N): communicator tests, bench at N, all-reduce sweep.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N"
nvidia-smi topo -m > gpurun_out/topo_$N.txt 2>&1
echo "== multi-gpu tests"
timeout 600 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider > gpurun_out/pytest_multi_$N.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_multi_$N.log | cut -c1-400
echo "== bench ours N=$N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 1000 --warmup 50 > gpurun_out/bench_ours_$N.json 2> gpurun_out/bench_ours_$N.err; echo "rc=$?"; cat gpurun_out/bench_ours_$N.json | tail -2; tail -5 gpurun_out/bench_ours_$N.err
echo "== bench reference N=$N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus $N --steps 1000 --warmup 50 > gpurun_out/bench_ref_$N.json 2> gpurun_out/bench_ref_$N.err; echo "rc=$?"; cat gpurun_out/bench_ref_$N.json | tail -2; tail -5 gpurun_out/bench_ref_$N.err
echo "== allreduce sweep N=$N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 benchmarks/allreduce_sweep.py --iters 100 --out gpurun_out/allreduce_$N.json > gpurun_out/allreduce_$N.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/allreduce_$N.log | cut -c1-300
