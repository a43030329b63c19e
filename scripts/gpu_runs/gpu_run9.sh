#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== engine tests"
timeout 300 python -u -m pytest tests/test_gpu_engine.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_engine.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_engine.log | cut -c1-300
echo "== e2e probe"
PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py 2>&1 | grep -v "^$" | grep -v "over 50 steps" | tee gpurun_out/e2e_probe.log
PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py --workers 10 --slots 16 2>&1 | grep -v "^$" | grep -v "over 50 steps" | tee -a gpurun_out/e2e_probe.log
echo "== bench"
timeout 300 python -u bench.py --steps 2000 --warmup 100 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "rc=$?"; cat gpurun_out/bench_ours.json; tail -3 gpurun_out/bench_ours.err
echo "== ncu (step kernel after the phase-2 / barrier / PDL changes)"
timeout 120 python benchmarks/profile_step.py --steps 12 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:net_kernel -s 6 -c 3 -f -o gpurun_out/prof_net_step_v2 python benchmarks/profile_step.py --steps 12 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_full.log
