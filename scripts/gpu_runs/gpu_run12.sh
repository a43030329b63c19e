#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 200 python -u benchmarks/cnn_breakdown.py --width 1 --batch 512 2>&1 | tee gpurun_out/cnn_breakdown_w1.log
timeout 300 python -u benchmarks/cnn_breakdown.py --width 8 --batch 128 2>&1 | tee gpurun_out/cnn_breakdown_w8.log
