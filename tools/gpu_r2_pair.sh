#!/bin/bash
# packed-pair explicit kernels: GPU parity tests of the default build, then same-box A/B of the variants
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log; tail -5 gpurun_out/r2g_pytest.log
for so in ${VARIANTS:-variants/base.so variants/scalar.so variants/pair_nopf.so}; do for w in c4_explicit c3_explicit; do
  FEMB200_LIB=$PWD/$so timeout 300 python bench.py --no-cpu --workload $w --steps 500 --warmup 10 --e2e-steps 3 2> gpurun_out/ab.err | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('$so $w', 'ms/step %.5f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'tile_ms', r.get('tile_ms'), 'bnd_ms', r.get('boundary_ms'), 'value %.4g'%d['value'])
except Exception as e: print('$so $w FAILED', e)
" | tee -a gpurun_out/r2g_ab.log
done; done
