#!/bin/bash
# Round-end check on one GPU: all parity tests, smoke(), the default bench line and the solver workloads.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 400 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; tail -c 1800 gpurun_out/bench_default.json
for w in c4_cg c4_jacobi; do
  timeout 300 python bench.py --workload $w --steps 200 --warmup 10 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; tail -c 700 gpurun_out/bench_$w.json
done
true
