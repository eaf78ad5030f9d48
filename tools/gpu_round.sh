#!/bin/bash
# One GPU visit: parity tests, the four bench workloads, launch lists.  Output under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
for w in c4_explicit c3_explicit; do
  python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 1500 gpurun_out/bench_$w.json
done
for w in c4_cg c4_jacobi; do
  python bench.py --workload $w --steps 200 --warmup 10 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 1200 gpurun_out/bench_$w.json
done
python bench.py --impl reference --steps 5 --warmup 4 > gpurun_out/bench_ref.json 2>&1; tail -c 600 gpurun_out/bench_ref.json
