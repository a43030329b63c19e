#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== tests"
timeout 400 python -u -m pytest tests/test_gpu_net_step.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_step.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_step.log | cut -c1-300
echo "== e2e probe"
PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py 2>&1 | grep -v "^$" | tee gpurun_out/e2e_probe.log
PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py --workers 10 --slots 16 2>&1 | grep -v "^$" | tee -a gpurun_out/e2e_probe.log
PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py --batch 64 2>&1 | grep -v "^$" | tee -a gpurun_out/e2e_probe.log
echo "== sweep"
timeout 120 python -u benchmarks/step_sweep.py --steps 3000 --batches 64,512,2048,8192 --ctas 0 2>&1 | grep batch | tee gpurun_out/step_sweep.log
PDM_NET_RT=8 timeout 120 python -u benchmarks/step_sweep.py --steps 3000 --batches 2048,8192 --ctas 0 2>&1 | grep batch | tee -a gpurun_out/step_sweep.log
PDM_NET_RT=4 timeout 120 python -u benchmarks/step_sweep.py --steps 3000 --batches 2048,8192 --ctas 0 2>&1 | grep batch | tee -a gpurun_out/step_sweep.log
echo "== bench"
timeout 300 python -u bench.py --steps 2000 --warmup 100 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "rc=$?"; cat gpurun_out/bench_ours.json; tail -3 gpurun_out/bench_ours.err
