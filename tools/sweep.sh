#!/bin/bash
# usage: tools/sweep.sh [bench args...]   -- runs bench.py with every variants/*.so
for so in variants/*.so; do
  FEMB200_LIB=$PWD/$so python bench.py --no-cpu --steps 300 --warmup 10 --e2e-steps 3 "$@" 2>&1 | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('$so', 'ms/step %.4f'%d['ms_per_step'], 'tile %.4f bnd %.4f'%(r['tile_ms'], r['boundary_ms']), 'tiles', d['config']['n_tiles_rank0'], 'Gtet/s %.1f'%(d['value']/1e9))
except Exception as e: print('$so FAILED', e)
"
done
