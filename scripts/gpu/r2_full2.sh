#!/usr/bin/env bash
# 2 GPUs: the whole GPU suite (single- and multi-GPU tests), log archived.
set -u
mkdir -p gpurun_out
export PYTHONFAULTHANDLER=1
timeout 1100 python -u -m pytest tests -m gpu -q -x --timeout=240 -p no:cacheprovider 2>&1 | tail -n 25 | cut -c1-400 | tee gpurun_out/pytest_gpu_2gpus.log
