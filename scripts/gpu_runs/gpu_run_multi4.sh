#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N  cores: $(nproc)"
echo "== single-GPU engine/step tests"
timeout 300 python -u -m pytest tests/test_gpu_engine.py tests/test_gpu_net_step.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_engine.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_engine.log | cut -c1-300
echo "== multi tests"
timeout 400 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider > gpurun_out/pytest_multi_$N.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_multi_$N.log | cut -c1-300
for n in 1 $N; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2961$n bench.py --gpus $n --steps 2000 --warmup 100 > gpurun_out/bench_ours_$n.json 2> gpurun_out/bench_ours_$n.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_ours_$n.json').read().strip().split('\n')[-1])
    print('ours N=%d value %.4g img/s (%.2f us/step)  e2e %.4g img/s (%.2f us/step, %s)' % (d['n_gpus'], d['value'], d['ms_per_step']*1e3, d['e2e']['value'], d['e2e']['ms_per_step']*1e3, d['e2e']['input_path'][:40]))
except Exception as e: print('parse error', e)
PY
done
PDM_NET_TRACE=1 timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 benchmarks/step_trace_multi.py 2>&1 | grep -v "OMP_NUM\|\*\*\*" | tee gpurun_out/step_trace_multi$N.log
