#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
echo "== net_step + engine + zz tests (default launch mode)"
timeout 300 python -u -m pytest tests/test_gpu_net_step.py tests/test_gpu_engine.py tests/test_zz_no_alloc.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_step.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_step.log | cut -c1-300
echo "== same tests, PDL launch, RT=4"
PDM_NET_LAUNCH=pdl PDM_NET_RT=4 timeout 300 python -u -m pytest tests/test_gpu_net_step.py tests/test_gpu_engine.py -m gpu -x -q --timeout=100 -p no:cacheprovider > gpurun_out/pytest_step_pdl.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_step_pdl.log | cut -c1-300
echo "== sweeps"
for mode in coop pdl; do for rt in 8 4; do
  echo "-- launch=$mode rt=$rt"
  PDM_NET_LAUNCH=$mode PDM_NET_RT=$rt timeout 120 python -u benchmarks/step_sweep.py --steps 3000 --batches 64,512 --ctas 0,64,128 2>&1 | grep batch
done; done | tee gpurun_out/step_sweep_modes.log
echo "== trace (coop rt8 / coop rt4)"
PDM_NET_TRACE=1 timeout 100 python -u benchmarks/step_trace.py --batch 512 > gpurun_out/step_trace2.log 2>&1
PDM_NET_TRACE=1 PDM_NET_RT=4 timeout 100 python -u benchmarks/step_trace.py --batch 512 >> gpurun_out/step_trace2.log 2>&1
cat gpurun_out/step_trace2.log
echo "== h2d probe"
timeout 100 python -u benchmarks/h2d_probe.py > gpurun_out/h2d_probe.log 2>&1; cat gpurun_out/h2d_probe.log
