#!/usr/bin/env bash
# 8 GPUs: phase trace of the persistent kernel's exchange at 4 and 8 ranks.
set -u
mkdir -p gpurun_out
for n in 4 8; do
  echo "== $n ranks"
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29570 + n)) \
    benchmarks/tc_trace_multi.py 2>&1 | grep -v -i "warn" | tail -n 16 | cut -c1-200
done | tee gpurun_out/tc_trace_multi.log
