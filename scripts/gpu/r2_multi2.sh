#!/usr/bin/env bash
# N GPUs: NVLS reducer tests, soak under skew, then the whole multi-GPU suite.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
timeout 300 python -u -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout=120 -p no:cacheprovider -k "in_the_switch" 2>&1 | tail -n 15 | cut -c1-300
echo "== soak (skew 20 us)"
PDM_SKEW_NS=20000 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29566 benchmarks/soak.py --steps ${SOAK_STEPS:-4000} --out gpurun_out/soak_${N}.json 2>&1 | grep -v Warning | tail -n 4 | cut -c1-900
