#!/usr/bin/env bash
# First GPU call of the next round (gpurun --gpus 2): everything that was written after the round-1
# GPU budget ran out, cheapest checks first.  Each step prints its own verdict; nothing here is timed.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
N=$(nvidia-smi -L | wc -l)
echo "== topology discovery on real hardware"
timeout 60 python -m pytorch_distributed_mnist_b200.parallel.topology 2>&1 | tail -14
echo "== multi-GPU suite (consolidation asserts, native CNN step, dead peer)"
#timeout 600 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=150 -p no:cacheprovider > gpurun_out/pytest_multi_r2_$N.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_multi_r2_$N.log | cut -c1-300
echo "== NVLS (experimental)"
PDM_TEST_NVLS=1 timeout 200 python -u -m pytest tests/test_gpu_multi.py -m gpu -x -q --timeout=120 -p no:cacheprovider -k nvls > gpurun_out/pytest_nvls_$N.log 2>&1; echo "rc=$?"; tail -25 gpurun_out/pytest_nvls_$N.log | cut -c1-300
echo "== all-reduce sweep with --nvls"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29613 benchmarks/allreduce_sweep.py --iters 50 --nvls --out gpurun_out/allreduce_nvls_$N.json 2>&1 | grep -v Warning | tail -14 | cut -c1-400
echo "== CLI job with --print-topology --comm-timeout (cnn, checkpoint with consolidated state)"
timeout 300 python -u -m pytorch_distributed_mnist_b200 --world-size $N --synthetic --train-size $((2048*N)) --test-size 1024 --batch-size $((64*N)) --epochs 1 --arch cnn --print-topology --comm-timeout 10 -i tcp://127.0.0.1:29655 --checkpoint-dir /tmp/ck_r2 2>&1 | grep -v Warning | tail -16 | cut -c1-200
