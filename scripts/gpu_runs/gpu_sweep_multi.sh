#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
run() { env "$@" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29681 benchmarks/step_sweep_multi.py 2>&1 | grep "us/step"; }
{
run PDM_NET_DEP=grid
run PDM_NET_DEP=flag
run PDM_NET_DEP=grid PDM_NET_LL=pull
run PDM_NET_DEP=flag PDM_NET_LL=pull
run PDM_NET_DEP=grid PDM_NET_LL=pull PDM_NET_RT=8
} | tee gpurun_out/step_sweep_multi_ll_$N.log
