#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
run() { env "$@" timeout 100 python -u benchmarks/step_sweep.py --steps 4000 --batches 512 --ctas 0 2>&1 | grep batch; }
{
run PDM_NET_DEP=grid PDM_NET_RT=4
run PDM_NET_DEP=flag PDM_NET_TRIGGER=barrier PDM_NET_RT=4
run PDM_NET_DEP=flag PDM_NET_TRIGGER=barrier PDM_NET_RT=4 PDM_NET_POLL_NS=100
run PDM_NET_DEP=flag PDM_NET_TRIGGER=start PDM_NET_RT=4
run PDM_NET_DEP=flag PDM_NET_TRIGGER=start PDM_NET_RT=4 PDM_NET_POLL_NS=100
run PDM_NET_DEP=flag PDM_NET_TRIGGER=start PDM_NET_RT=8
run PDM_NET_DEP=flag PDM_NET_TRIGGER=start PDM_NET_RT=8 PDM_NET_POLL_NS=100
run PDM_NET_DEP=flag PDM_NET_TRIGGER=start PDM_NET_RT=8 PDM_NET_POLL_NS=500
run PDM_NET_DEP=grid PDM_NET_RT=8
} | tee gpurun_out/step_sweep_dep.log
echo "== correctness under the fastest candidates"
PDM_NET_DEP=flag PDM_NET_TRIGGER=start timeout 200 python -u -m pytest tests/test_gpu_engine.py -m gpu -x -q --timeout=100 -p no:cacheprovider 2>&1 | tail -2
