#!/usr/bin/env bash
# Round-end single-GPU validation: smoke, full GPU suite, both bench arms, memcheck.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== smoke"; timeout 200 python __graft_entry__.py smoke 2>&1 | tail -1
echo "== full gpu suite"
timeout 600 python -u -m pytest tests -m gpu -x -q --timeout=120 -p no:cacheprovider > gpurun_out/pytest_gpu_all.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_gpu_all.log | cut -c1-300
echo "== bench reference"
timeout 600 python -u bench.py --impl reference --steps 2000 --warmup 100 > gpurun_out/bench_ref_1.json 2> gpurun_out/bench_ref_1.err; echo "rc=$?"
echo "== bench ours"
timeout 300 python -u bench.py --steps 2000 --warmup 100 > gpurun_out/bench_ours_1.json 2> gpurun_out/bench_ours_1.err; echo "rc=$?"
python - <<PY
import json
for arm in ('ref','ours'):
    try:
        d=json.loads(open('gpurun_out/bench_%s_1.json'%arm).read().strip().split('\n')[-1])
        print(arm, 'value %.4g img/s (%.2f us/step)  e2e %.4g img/s (%.2f us/step) clocks %s launches %s' % (d['value'], d['ms_per_step']*1e3, d['e2e']['value'], d['e2e']['ms_per_step']*1e3, d['clocks'], d.get('gpu_launches')))
    except Exception as e: print(arm, 'parse error', e)
PY
echo "== compute-sanitizer memcheck"
timeout 200 python scripts/sanitize_smoke.py > gpurun_out/sanitize_plain.log 2>&1 && \
timeout 900 compute-sanitizer --tool memcheck --log-file gpurun_out/memcheck.log python scripts/sanitize_smoke.py > gpurun_out/sanitize_run.log 2>&1
echo "sanitizer rc=$?"; tail -3 gpurun_out/sanitize_run.log; grep -E "ERROR SUMMARY|Invalid|error" gpurun_out/memcheck.log | head -8
