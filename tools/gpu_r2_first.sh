#!/bin/bash
# Round 2, first visit: full GPU suite on the default build and on the persistent-explicit variant, then a same-box A/B.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2_first_smi.txt
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_default.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_default.log; tail -3 gpurun_out/r2_pytest_default.log
FEMB200_LIB=$PWD/variants/persist_explicit.so timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_pe.log 2>&1; echo "rc=$?" >> gpurun_out/r2_pytest_pe.log; tail -3 gpurun_out/r2_pytest_pe.log
for rep in 1 2 3; do for so in base persist_explicit; do
  FEMB200_LIB=$PWD/variants/$so.so timeout 300 python bench.py --no-cpu --workload c4_explicit --steps 500 --warmup 10 --e2e-steps 3 2>> gpurun_out/r2_ab.err | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$so rep$rep', 'ms/step %.4f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'tile_ms', r.get('tile_ms'), 'bnd_ms', r.get('boundary_ms'))
" | tee -a gpurun_out/r2_ab.log
done; done
