# Convenience targets; everything also works as the plain commands shown in README.md.
PY ?= python

.PHONY: build test test-gpu bench bench-ref sass clean

build:            ## compile the native library for sm_100a (nvcc; no GPU needed)
	$(PY) -m pytorch_distributed_mnist_b200.ops._build

test: build       ## CPU suite (gloo, world_size 2-3)
	$(PY) -m pytest tests -x -q -m "not gpu"

test-gpu: build   ## GPU suite (multi-GPU tests run when >= 2 devices are visible)
	$(PY) -m pytest tests -x -q -m gpu

bench: build      ## headline benchmark, 1 GPU (N>1: see bench.py docstring)
	$(PY) bench.py --gpus 1

bench-ref:        ## the unmodified reference through the same harness
	$(PY) bench.py --gpus 1 --impl reference

sass: build       ## regenerate profiles/sass/
	$(PY) scripts/sass_summary.py

clean:
	rm -rf build pytorch_distributed_mnist_b200/_native
