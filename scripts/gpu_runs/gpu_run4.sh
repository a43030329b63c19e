#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== no-alloc test alone (fresh process)"
timeout 200 python -u -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_zz_alone.log | cut -c1-300
echo "== full gpu suite (-x)"
timeout 600 python -u -m pytest tests -m gpu -x -q --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_gpu_all.log | cut -c1-300
echo "== step sweep"
timeout 200 python -u benchmarks/step_sweep.py --steps 2000 --batches 64,512,4096 --ctas 0,32,64,128 > gpurun_out/step_sweep.log 2>&1; echo "rc=$?"; cat gpurun_out/step_sweep.log
echo "== bench ours"
timeout 300 python -u bench.py --steps 2000 --warmup 100 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "rc=$?"; cat gpurun_out/bench_ours.json; tail -5 gpurun_out/bench_ours.err
