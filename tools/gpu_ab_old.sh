#!/bin/bash
# same-box A/B of the current library against one built from an older commit (variants/old_<commit>.so)
mkdir -p gpurun_out
for so in variants/head.so variants/old_*.so variants/head.so; do for w in c4_explicit c4_cg c4_jacobi; do
  S=200; [[ $w == *explicit ]] && S=500
  FEMB200_ALLOW_OLD_LIB=1 FEMB200_LIB=$PWD/$so timeout 300 python bench.py --no-cpu --no-parity --workload $w --steps $S --warmup 10 --e2e-steps 3 2> gpurun_out/abo.err | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('$so $w', 'ms/step %.5f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'tile_ms', r.get('tile_ms'), 'bnd_ms', r.get('boundary_ms'))
except Exception as e: print('$so $w FAILED', e)
" | tee -a gpurun_out/r2_ab_old.log
done; done
tail -5 gpurun_out/abo.err
