#!/bin/bash
# final one-GPU visit of round 2: same-box A/B against the round-start build, GPU parity tests, smoke, the default
# bench line + reference arm, ncu launch lists and ncu --set full of the dominant kernels
mkdir -p gpurun_out
bash tools/gpu_ab_old.sh
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2x_pytest.log; tail -3 gpurun_out/r2x_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2x_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r2x_smoke.log
python bench.py > gpurun_out/r2x_bench_n1.json 2> gpurun_out/r2x_bench_n1.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2x_bench_ref.json 2> gpurun_out/r2x_bench_ref.err; echo "ref rc=$?"
B="python bench.py --no-cpu --no-parity --e2e-steps 3"
for w in c4_explicit c4_cg c4_jacobi; do
  $B --workload $w --steps 10 --warmup 3 > gpurun_out/plain_$w.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2x_launches_$w.csv $B --workload $w --steps 10 --warmup 3 > gpurun_out/ncu_$w.log 2>&1
done
ncu --set full --clock-control none --import-source on -k regex:k_tile -s 6 -c 1 -o gpurun_out/r2x_prof_explicit $B --workload c4_explicit --steps 10 --warmup 3 > gpurun_out/ncu_full_ex.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile_persistent -s 8 -c 2 -o gpurun_out/r2x_prof_apply $B --workload c4_cg --steps 10 --warmup 3 > gpurun_out/ncu_full_cg.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_tile_persistent -s 6 -c 1 -o gpurun_out/r2x_prof_jacobi $B --workload c4_jacobi --steps 10 --warmup 3 > gpurun_out/ncu_full_jac.log 2>&1
ls -la gpurun_out/r2x*.ncu-rep
