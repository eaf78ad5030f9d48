#!/bin/bash
# after the ITC code left the default build: same-box A/B against older builds, the default bench line, ncu of the
# operator / explicit kernels, launch lists
mkdir -p gpurun_out
bash tools/gpu_ab_old.sh
python bench.py > gpurun_out/r2y_bench_n1.json 2> gpurun_out/r2y_bench_n1.err; echo "bench rc=$?"
B="python bench.py --no-cpu --no-parity --e2e-steps 3"
for w in c4_explicit c4_cg c4_jacobi; do
  $B --workload $w --steps 10 --warmup 3 > gpurun_out/plain_$w.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2y_launches_$w.csv $B --workload $w --steps 10 --warmup 3 > gpurun_out/ncu_$w.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:k_tile -s 6 -c 1 -o gpurun_out/r2y_prof_explicit $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_full_ex.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile_persistent -s 8 -c 2 -o gpurun_out/r2y_prof_apply $B --workload c4_cg --steps 10 --warmup 3 > gpurun_out/ncu_full_cg.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile_persistent -s 6 -c 1 -o gpurun_out/r2y_prof_jacobi $B --workload c4_jacobi --steps 10 --warmup 3 > gpurun_out/ncu_full_jac.log 2>&1
ls -la gpurun_out/r2y*.ncu-rep
