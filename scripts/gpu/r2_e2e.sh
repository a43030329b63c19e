#!/usr/bin/env bash
set -u
export PYTHONFAULTHANDLER=1
timeout 200 python -u -m pytest tests/test_gpu_net_tc.py tests/test_gpu_engine.py -m gpu -q -x --timeout=60 -p no:cacheprovider 2>&1 | tail -n 12 | cut -c1-300
for k in 20 2000; do echo "== zerocopy steps=$k"; timeout 200 python bench.py --gpus 1 --steps $k --warmup 5 --placement zerocopy --no-cnn 2>&1 | grep -v Warning | tail -n 1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d[\"ms_per_step\"]*1e3,2), \"us/step value;\", \"e2e wall\", round(d[\"e2e\"][\"ms_per_step\"]*1e3,2), \"dev\", round(d[\"e2e\"][\"device_ms_per_step\"]*1e3,2), d[\"gpu_launches\"])"; done
