#!/bin/bash
# One-GPU validation + profiling visit: parity tests, smoke, the four bench workloads, the reference arm,
# ncu launch lists of the bench commands and ncu --set full of the dominant kernels.  Output: gpurun_out/
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
python bench.py > gpurun_out/bench_c4_explicit.json 2> gpurun_out/bench_c4_explicit.err; tail -c 300 gpurun_out/bench_c4_explicit.json; echo
for w in c3_explicit c4_cg c4_jacobi; do
  S=200; [ $w = c3_explicit ] && S=1000
  python bench.py --workload $w --steps $S --warmup 10 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
done
python bench.py --impl reference --steps 200 --warmup 5 > gpurun_out/bench_ref.json 2>&1
B="python bench.py --no-cpu --e2e-steps 3"
$B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/plain_ex.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_explicit.csv $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_ex.log 2>&1
$B --workload c4_cg --steps 20 --warmup 3 > gpurun_out/plain_cg.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_cg.csv $B --workload c4_cg --steps 20 --warmup 3 > gpurun_out/ncu_cg.log 2>&1
$B --workload c4_jacobi --steps 20 --warmup 3 > gpurun_out/plain_jac.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_jacobi.csv $B --workload c4_jacobi --steps 20 --warmup 3 > gpurun_out/ncu_jac.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile -s 6 -c 1 -o gpurun_out/prof_explicit $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_full_ex.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile_persistent -s 8 -c 2 -o gpurun_out/prof_apply $B --workload c4_cg --steps 20 --warmup 3 > gpurun_out/ncu_full_cg.log 2>&1
ncu --set full --clock-control none -k regex:k_boundary -s 6 -c 1 -o gpurun_out/prof_boundary $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_full_bnd.log 2>&1
ls -la gpurun_out/*.ncu-rep
