#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== debug engine"
PDM_ENGINE_DEBUG=1 timeout 120 python -u scripts/debug_engine.py > gpurun_out/debug_engine.log 2>&1; echo "rc=$?"; grep debug_engine gpurun_out/debug_engine.log | tail -8
for f in test_gpu_gemm test_gpu_net_step test_gpu_engine test_gpu_cnn test_gpu_ops; do
  echo "== pytest $f"
  timeout 300 python -u -m pytest tests/$f.py -m gpu -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_$f.log 2>&1; echo "rc=$?"
  tail -6 gpurun_out/pytest_$f.log | cut -c1-300
  grep -E "^(FAILED|ERROR)" gpurun_out/pytest_$f.log | cut -c1-200 | head -20
done
echo "== step sweep"
timeout 200 python -u benchmarks/step_sweep.py --steps 2000 > gpurun_out/step_sweep.log 2>&1; echo "rc=$?"; cat gpurun_out/step_sweep.log
PDM_NET_COOP=0 timeout 200 python -u benchmarks/step_sweep.py --steps 2000 --batches 512 --ctas 0,32 > gpurun_out/step_sweep_nocoop.log 2>&1; echo "rc=$?"; cat gpurun_out/step_sweep_nocoop.log
echo "== bench ours"
timeout 300 python -u bench.py --steps 2000 --warmup 100 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "rc=$?"; cat gpurun_out/bench_ours.json; tail -5 gpurun_out/bench_ours.err
