#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 120 python -u -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout=100 -p no:cacheprovider -k "persistent_tc" 2>&1 | tail -n 3 | cut -c1-200
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $N --steps 200 --warmup 20 --no-cnn --no-allreduce 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_${N}_ll2.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], 'GPUs:', round(d['ms_per_step']*1e3,2), 'us/step', round(d['value']/1e6,1), 'M img/s; fp32-input (per-step kernel)', round(d['fp32_inputs']['ms_per_step']*1e3,2), '; e2e wall', round(d['e2e']['ms_per_step']*1e3,2), 'dev', round(d['e2e']['device_ms_per_step']*1e3,2))"
