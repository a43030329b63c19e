#!/usr/bin/env bash
# 1 GPU: ncu --set full of the CNN step's FMA kernels and the bucket kernel (first step of a 2-step run).
set -u
mkdir -p gpurun_out
timeout 50 python benchmarks/profile_cnn_step.py 512 2 2>&1 | tail -n 1 || exit 1
timeout 85 ncu --set full --clock-control none --import-source on -k regex:'conv1_fwd|conv1_bwd|pool_fwd|pool_bwd|fcout_fwd|fcout_bwd|reduce_optim' -c 8 -f -o gpurun_out/prof_cnn_fma python benchmarks/profile_cnn_step.py 512 1 2>&1 | tail -n 3 | cut -c1-200
ls -la gpurun_out/*.ncu-rep 2>/dev/null
