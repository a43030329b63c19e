#!/usr/bin/env bash
# N GPUs: the headline bench with the persistent kernel forced on and off for resident inputs.
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
for mw in 8 0; do
echo "== PDM_NET_TC_MAX_WORLD=$mw"
PDM_NET_TC_MAX_WORLD=$mw timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2970$mw bench.py --gpus $N --steps 200 --warmup 20 --no-cnn --no-allreduce 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_${N}_maxworld${mw}.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], 'GPUs:', round(d['ms_per_step']*1e3,2), 'us/step', round(d['value']/1e6,1), 'M img/s launches', d['gpu_launches'], '; fp32-input (per-step kernel)', round(d['fp32_inputs']['ms_per_step']*1e3,2), '; e2e wall', round(d['e2e']['ms_per_step']*1e3,2), 'dev', round(d['e2e']['device_ms_per_step']*1e3,2))"
done
