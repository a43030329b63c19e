#!/usr/bin/env bash
# Second GPU pass: find the engine hang, validate GEMM/CNN kernels.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== debug engine"
PDM_ENGINE_DEBUG=1 timeout 150 python -u scripts/debug_engine.py > gpurun_out/debug_engine.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/debug_engine.log
for f in test_gpu_net_step test_gpu_ops test_gpu_gemm test_gpu_cnn test_gpu_engine; do
  echo "== pytest $f"
  timeout 400 python -u -m pytest tests/$f.py -m gpu -q --timeout=120 -p no:cacheprovider -x > gpurun_out/pytest_$f.log 2>&1; echo "rc=$?"
  tail -12 gpurun_out/pytest_$f.log
done
echo "== bench ours (short)"
timeout 300 python -u bench.py --steps 200 --warmup 10 > gpurun_out/bench_ours_short.json 2> gpurun_out/bench_ours_short.err; echo "rc=$?"; cat gpurun_out/bench_ours_short.json; tail -5 gpurun_out/bench_ours_short.err
