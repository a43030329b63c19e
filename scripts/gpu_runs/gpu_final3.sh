#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
timeout 120 python benchmarks/profile_gemm.py > gpurun_out/gemm_plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tf32 -s 3 -c 2 -f -o gpurun_out/prof_gemm python benchmarks/profile_gemm.py > gpurun_out/ncu_gemm.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/gemm_plain.log; tail -2 gpurun_out/ncu_gemm.log
