#!/bin/bash
# A/B of library variants (variants/*.so) on the bench workloads, after the GPU parity tests of the default build
mkdir -p gpurun_out
[ -n "$SKIP_TESTS" ] || timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
for so in variants/*.so; do for w in c4_explicit c4_cg c4_jacobi; do
  S=200; [[ $w == *explicit ]] && S=500
  FEMB200_LIB=$PWD/$so timeout 300 python bench.py --no-cpu --workload $w --steps $S --warmup 10 --e2e-steps 3 2> gpurun_out/ab.err | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('$so $w', 'ms/step %.4f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'tile_ms', r.get('tile_ms'), 'bnd_ms', r.get('boundary_ms'), 'value %.4g'%d['value'])
except Exception as e: print('$so $w FAILED', e)
" | tee -a gpurun_out/ab.log
done; done
