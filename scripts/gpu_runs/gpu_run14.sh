#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
echo "== engine + step tests"
timeout 300 python -u -m pytest tests/test_gpu_engine.py tests/test_gpu_net_step.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_engine.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/pytest_engine.log | cut -c1-300
echo "== e2e probe by placement"
for pl in staged device zerocopy; do
  timeout 120 python -u benchmarks/e2e_probe.py --placement $pl 2>&1 | grep -E "shuffled|sequential"
done | tee gpurun_out/e2e_placement.log
timeout 120 python -u benchmarks/e2e_probe.py --placement zerocopy --batch 64 2>&1 | grep -E "shuffled" | tee -a gpurun_out/e2e_placement.log
timeout 120 python -u benchmarks/e2e_probe.py --placement device --batch 64 2>&1 | grep -E "shuffled" | tee -a gpurun_out/e2e_placement.log
