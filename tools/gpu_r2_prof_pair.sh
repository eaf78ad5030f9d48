#!/bin/bash
# rest of the GPU tests + ncu --set full of the packed explicit tile kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_fullsize.py > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log; tail -5 gpurun_out/r2h_pytest.log
B="python bench.py --no-cpu --e2e-steps 3 --no-parity"
$B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/plain_ex.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_tile -s 6 -c 1 -o gpurun_out/prof_explicit_pair $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_full_ex.log 2>&1
ls -la gpurun_out/*.ncu-rep
