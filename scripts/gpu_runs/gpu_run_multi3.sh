#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N  cores: $(nproc)"
timeout 400 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider -k "ll_allreduce or full_job" > gpurun_out/pytest_multi_$N.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_multi_$N.log | cut -c1-300
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 2000 --warmup 100 > gpurun_out/bench_ours_$N.json 2> gpurun_out/bench_ours_$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_ours_$N.json').read().strip().split('\n')[-1])
    print('ours N=%d value %.4g img/s (%.2f us/step)  e2e %.4g img/s (%.2f us/step, %s)' % (d['n_gpus'], d['value'], d['ms_per_step']*1e3, d['e2e']['value'], d['e2e']['ms_per_step']*1e3, d['e2e']['input_path'][:40]))
except Exception as e: print('parse error', e)
PY
