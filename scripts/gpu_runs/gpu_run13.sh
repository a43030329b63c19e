#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
echo "== gemm + cnn tests"
timeout 300 python -u -m pytest tests/test_gpu_gemm.py tests/test_gpu_cnn.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_cnn.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/pytest_cnn.log | cut -c1-300
timeout 200 python -u benchmarks/cnn_breakdown.py --width 1 --batch 512 2>&1 | tee gpurun_out/cnn_breakdown_w1.log
timeout 300 python -u benchmarks/cnn_breakdown.py --width 8 --batch 128 2>&1 | tee gpurun_out/cnn_breakdown_w8.log
