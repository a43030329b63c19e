#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== no-alloc test alone (fresh process)"
timeout 200 python -u -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_zz_alone.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_zz_alone.log | cut -c1-300
echo "== full gpu suite (-x)"
timeout 600 python -u -m pytest tests -m gpu -x -q --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/pytest_gpu_all.log | cut -c1-300
echo "== step trace"
for b in 64 512 4096; do PDM_NET_TRACE=1 timeout 100 python -u benchmarks/step_trace.py --batch $b; done > gpurun_out/step_trace.log 2>&1; cat gpurun_out/step_trace.log
PDM_NET_TRACE=1 timeout 100 python -u benchmarks/step_trace.py --batch 512 --ctas 128 >> gpurun_out/step_trace.log 2>&1; tail -7 gpurun_out/step_trace.log
