#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
echo "== cnn tests"
timeout 300 python -u -m pytest tests/test_gpu_cnn.py tests/test_gpu_gemm.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_cnn.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_cnn.log | cut -c1-300
echo "== cnn bench width 1"
timeout 300 python -u benchmarks/bench_cnn.py --width 1 --batch 512 --steps 50 --warmup 10 --out gpurun_out/cnn_w1_1.json 2>&1 | grep -E "impl|Error|error" | cut -c1-400
echo "== cnn bench width 8"
timeout 400 python -u benchmarks/bench_cnn.py --width 8 --batch 128 --steps 20 --warmup 5 --out gpurun_out/cnn_w8_1.json 2>&1 | grep -E "impl|Error|error" | cut -c1-400
