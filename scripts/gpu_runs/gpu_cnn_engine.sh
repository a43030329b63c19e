#!/usr/bin/env bash
# Validate and time the native CNN step executor on one GPU.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== cnn tests"
timeout 400 python -u -m pytest tests/test_gpu_cnn.py -m gpu -x -q --timeout=120 -p no:cacheprovider > gpurun_out/pytest_cnn_engine.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_cnn_engine.log | cut -c1-400
echo "== bench w1"
timeout 200 python -u benchmarks/bench_cnn.py --width 1 --batch 512 --steps 100 --warmup 20 --out gpurun_out/cnn_engine_w1.json 2>&1 | tail -5
echo "== bench w8"
timeout 200 python -u benchmarks/bench_cnn.py --width 8 --batch 128 --steps 50 --warmup 10 --out gpurun_out/cnn_engine_w8.json 2>&1 | tail -5
