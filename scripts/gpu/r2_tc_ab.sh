#!/usr/bin/env bash
# 1 GPU: A/B of builds of the library on the same box (each build_alt/*.so replaces the in-tree one in turn).
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
run() {
  timeout 200 python benchmarks/tc_sweep.py --steps 2000 --batches 64,512 2>&1 | grep -v Warning | tail -n 2 | cut -c1-75
}
echo "== A: in-tree build"; run
cp pytorch_distributed_mnist_b200/_native/libpdmnist_b200.so /tmp/libA.so
for alt in build_alt/*.so; do
  [ -f "$alt" ] || continue
  cp "$alt" pytorch_distributed_mnist_b200/_native/libpdmnist_b200.so
  echo "== B: $alt"; run
done
cp /tmp/libA.so pytorch_distributed_mnist_b200/_native/libpdmnist_b200.so
