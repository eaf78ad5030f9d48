#!/bin/bash
# 2-GPU visit: the 2-rank slab check (both transports), then the scaling script
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -12 gpurun_out/pytest_gpu2.log | cut -c1-600
bash tools/gpu_scale.sh "$@"
