#!/usr/bin/env bash
# 1 GPU: tests of the tensor-core cluster kernel, launch-cost probe, batch sweep.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
timeout 400 python -u -m pytest tests/test_gpu_net_tc.py tests/test_gpu_engine.py -m gpu -q -x --timeout=120 -p no:cacheprovider 2>&1 | tail -n 12 | cut -c1-300
timeout 120 python benchmarks/launch_cost_probe.py 2>&1 | grep -v Warning | tail -n 10
timeout 200 python benchmarks/tc_sweep.py --steps 2000 2>&1 | grep -v Warning | tail -n 6 | tee gpurun_out/tc_sweep.jsonl
