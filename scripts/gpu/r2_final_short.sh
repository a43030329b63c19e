#!/usr/bin/env bash
# 1 GPU, short form of the round-end rehearsal: Net kernels + engine + allocation-free entry tests, smoke(), bench at 20 steps.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
timeout 70 python -u -m pytest tests/test_gpu_net_tc.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=60 -p no:cacheprovider 2>&1 | tail -n 3 | cut -c1-200
timeout 40 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -n 2 | cut -c1-300
timeout 60 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cnn --no-allreduce 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_1_final.json | cut -c1-1800
