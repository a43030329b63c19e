#!/usr/bin/env bash
# Multi-GPU pass v2: tests, bench (both arms), CNN bench width 8, all-reduce sweep.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
echo "GPUs: $N  cores: $(nproc)"
echo "== multi-gpu tests"
timeout 600 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider > gpurun_out/pytest_multi_$N.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/pytest_multi_$N.log | cut -c1-300
echo "== bench ours N=$N"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 2000 --warmup 100 > gpurun_out/bench_ours_$N.json 2> gpurun_out/bench_ours_$N.err; echo "rc=$?"; tail -3 gpurun_out/bench_ours_$N.err | cut -c1-300
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_ours_$N.json').read().strip().split('\n')[-1])
    print('ours N=%d value %.4g img/s (%.2f us/step)  e2e %.4g img/s (%.2f us/step, %s)' % (d['n_gpus'], d['value'], d['ms_per_step']*1e3, d['e2e']['value'], d['e2e']['ms_per_step']*1e3, d['e2e']['input_path'][:40]))
except Exception as e: print('parse error', e)
PY
echo "== bench ours N=$N staged e2e"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29614 bench.py --gpus $N --steps 1000 --warmup 50 --placement staged > gpurun_out/bench_ours_staged_$N.json 2> /dev/null; python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_ours_staged_$N.json').read().strip().split('\n')[-1])
    print('staged e2e %.4g img/s (%.2f us/step)' % (d['e2e']['value'], d['e2e']['ms_per_step']*1e3))
except Exception as e: print('parse error', e)
PY
echo "== bench reference N=$N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29612 bench.py --impl reference --gpus $N --steps 1000 --warmup 50 > gpurun_out/bench_ref_$N.json 2> gpurun_out/bench_ref_$N.err; echo "rc=$?"
python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_ref_$N.json').read().strip().split('\n')[-1])
    print('ref N=%d value %.4g img/s (%.2f us/step)  e2e %.4g img/s' % (d['n_gpus'], d['value'], d['ms_per_step']*1e3, d['e2e']['value']))
except Exception as e: print('parse error', e)
PY
echo "== cnn width 8 N=$N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29615 benchmarks/bench_cnn.py --width 8 --batch 128 --steps 20 --warmup 5 --out gpurun_out/cnn_w8_$N.json 2>&1 | grep -E "impl|Error" | cut -c1-300
echo "== cnn width 1 N=$N"
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29616 benchmarks/bench_cnn.py --width 1 --batch 512 --steps 30 --warmup 5 --out gpurun_out/cnn_w1_$N.json 2>&1 | grep -E "impl|Error" | cut -c1-300
