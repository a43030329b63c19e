#!/usr/bin/env bash
# 1 GPU: tests of the tensor-core cluster kernel + engine regression, phase trace, step sweep.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
timeout 600 python -u -m pytest tests/test_gpu_net_tc.py tests/test_gpu_engine.py -m gpu -q -x --timeout=120 -p no:cacheprovider 2>&1 | tail -n 25 | cut -c1-300
echo "== trace"
PDM_TC_TRACE=1 timeout 120 python benchmarks/tc_trace.py 512 2>&1 | grep -v Warning | tail -n 16 | tee gpurun_out/tc_trace_512.txt | cut -c1-230
echo "== sweep"
timeout 300 python benchmarks/tc_sweep.py --steps 2000 2>&1 | grep -v Warning | tail -n 6 | tee gpurun_out/tc_sweep.jsonl
