#!/usr/bin/env bash
# First GPU pass: smoke, GPU tests, both bench arms, launch list + one full ncu capture.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
echo "== smoke" ; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
echo "== bench ours"; timeout 600 python bench.py --steps 2000 --warmup 100 > gpurun_out/bench_ours.json 2> gpurun_out/bench_ours.err; echo "rc=$?"; cat gpurun_out/bench_ours.json; tail -3 gpurun_out/bench_ours.err
echo "== bench reference"; timeout 900 python bench.py --impl reference --steps 2000 --warmup 100 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "rc=$?"; cat gpurun_out/bench_ref.json; tail -3 gpurun_out/bench_ref.err
echo "== cta sweep"
for c in 16 32 64 128 148; do timeout 120 python benchmarks/profile_step.py --steps 200 --ctas $c; done > gpurun_out/cta_sweep.log 2>&1; cat gpurun_out/cta_sweep.log
echo "== ncu"
timeout 120 python benchmarks/profile_step.py --steps 12 > gpurun_out/plain.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:net_kernel -s 6 -c 3 -f -o gpurun_out/prof_net_step python benchmarks/profile_step.py --steps 12 > gpurun_out/ncu_full.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_full.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv python benchmarks/profile_step.py --steps 12 > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
