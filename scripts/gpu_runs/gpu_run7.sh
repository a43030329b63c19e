#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
for mode in coop pdl; do
  echo "-- launch=$mode"
  PDM_NET_LAUNCH=$mode PDM_NET_RT=4 PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py 2>&1 | grep -v "^$"
done | tee gpurun_out/e2e_probe.log
PDM_NET_LAUNCH=pdl PDM_NET_RT=4 PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py --workers 12 --slots 16 2>&1 | tee -a gpurun_out/e2e_probe.log
PDM_NET_LAUNCH=pdl PDM_NET_RT=4 PDM_ENGINE_PROFILE=1 timeout 120 python -u benchmarks/e2e_probe.py --batch 64 2>&1 | tee -a gpurun_out/e2e_probe.log
nproc; lscpu | grep -E "Model name|^CPU\(s\)|Socket|Thread" 
