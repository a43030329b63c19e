#!/usr/bin/env bash
# Multi-GPU validation of the native CNN step executor (gpurun --gpus N).
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N"
timeout 400 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider -k "cnn" > gpurun_out/pytest_cnn_multi_$N.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_cnn_multi_$N.log | cut -c1-400
echo "== cnn job through the CLI (Trainer -> CnnEngine -> peer reducer)"
timeout 300 python -u -m pytorch_distributed_mnist_b200 --world-size $N --synthetic --train-size $((4096*N)) --test-size 2048 --batch-size $((64*N)) --epochs 2 --arch cnn -i tcp://127.0.0.1:29655 --checkpoint-dir /tmp/ck_cnn > gpurun_out/cnn_job_$N.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/cnn_job_$N.log | cut -c1-300
for W in 1 8; do
  B=512; [ $W = 8 ] && B=128
  echo "== bench_cnn width $W N=$N"
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$W benchmarks/bench_cnn.py --width $W --batch $B --steps 50 --warmup 10 --out gpurun_out/cnn_engine_w${W}_$N.json 2>&1 | grep impl
done
