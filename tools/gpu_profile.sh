#!/bin/bash
# One-GPU profiling visit: launch lists + ncu --set full of the tile kernels (explicit step, implicit apply)
mkdir -p gpurun_out
B="python bench.py --no-cpu --e2e-steps 3"
$B --workload c4_cg --steps 20 --warmup 3 > gpurun_out/plain_cg.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cg.csv $B --workload c4_cg --steps 20 --warmup 3 > gpurun_out/ncu_cg.log 2>&1
$B --workload c4_jacobi --steps 20 --warmup 3 > gpurun_out/plain_jac.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_jacobi.csv $B --workload c4_jacobi --steps 20 --warmup 3 > gpurun_out/ncu_jac.log 2>&1
$B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/plain_ex.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_explicit.csv $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_ex.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile -s 12 -c 2 -o gpurun_out/prof_apply $B --workload c4_cg --steps 20 --warmup 3 > gpurun_out/ncu_full_cg.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile -s 6 -c 1 -o gpurun_out/prof_explicit $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_full_ex.log 2>&1
ls -la gpurun_out/*.ncu-rep gpurun_out/*.csv
