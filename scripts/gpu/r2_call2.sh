#!/usr/bin/env bash
# Round-2 call 2 (1 GPU): e2e wall-clock probe after moving set-up out of train(); quick regression of the touched paths.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
timeout 600 python -u -m pytest tests/test_gpu_engine.py tests/test_gpu_net_step.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=200 -p no:cacheprovider 2>&1 | tail -n 5
for pl in staged zerocopy; do for k in 20 2000; do
echo "== bench placement=$pl steps=$k"
PDM_ENGINE_PROFILE=1 timeout 300 python bench.py --gpus 1 --steps $k --warmup 5 --placement $pl --no-cnn 2>&1 | grep -v Warning | tail -n 4 | cut -c1-2400
done; done
echo "== default bench (with cnn extension)"
timeout 400 python bench.py --gpus 1 --steps 20 --warmup 5 2>&1 | grep -v Warning | tail -n 1 | tee gpurun_out/bench_ours_1_r2b.json | cut -c1-3000
