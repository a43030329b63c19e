#!/usr/bin/env bash
# N GPUs: the persistent-kernel multi-GPU test, then bench.py with the persistent kernel forced on and off.
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 200 python -u -m pytest tests/test_gpu_multi.py -m gpu -q -x --timeout=150 -p no:cacheprovider -k "persistent_tc or fused_net_step" 2>&1 | tail -n 3 | cut -c1-200
for mw in 8 0; do
  PDM_NET_TC_MAX_WORLD=$mw timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29701 + mw)) bench.py --gpus $N --steps ${STEPS:-20} --warmup 5 --no-cnn --no-allreduce 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_${N}_maxworld${mw}.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('max world $mw:', d['n_gpus'], 'GPUs:', round(d['ms_per_step']*1e3,2), 'us/step', round(d['value']/1e6,1), 'M img/s; launches', d.get('gpu_launches'), '; fp32-input (per-step kernel)', round(d['fp32_inputs']['ms_per_step']*1e3,2), '; e2e wall', round(d['e2e']['ms_per_step']*1e3,2), 'dev', round(d['e2e']['device_ms_per_step']*1e3,2))"
done
