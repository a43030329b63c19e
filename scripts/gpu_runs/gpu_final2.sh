#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
PDM_NET_TRACE=1 timeout 100 python -u benchmarks/step_trace.py --batch 512 2>&1 | tee gpurun_out/step_trace3.log
timeout 100 python -u benchmarks/step_sweep.py --steps 4000 --batches 512 --ctas 0 2>&1 | grep batch
echo "== compute-sanitizer memcheck"
timeout 200 python scripts/sanitize_smoke.py > gpurun_out/sanitize_plain.log 2>&1 && \
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/memcheck.log python scripts/sanitize_smoke.py > gpurun_out/sanitize_run.log 2>&1
echo "sanitizer rc=$?"; tail -2 gpurun_out/sanitize_plain.log; tail -3 gpurun_out/sanitize_run.log; grep -E "ERROR SUMMARY|Invalid|error" gpurun_out/memcheck.log | head -8
