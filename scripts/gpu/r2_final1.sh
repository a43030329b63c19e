#!/usr/bin/env bash
# 1 GPU, what the driver runs at round end: full GPU suite, smoke(), both bench arms at 20 steps.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
timeout 900 python -u -m pytest tests -m gpu -x -q --timeout=200 -p no:cacheprovider > gpurun_out/pytest_gpu_all_1gpu.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/pytest_gpu_all_1gpu.log | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2
timeout 300 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ref_1_r2.json | cut -c1-600
timeout 400 python bench.py --gpus 1 --steps 20 --warmup 5 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_1_r2.json | cut -c1-3000
