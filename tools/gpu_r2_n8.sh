#!/bin/bash
# 8-GPU visit: the default bench line with parity, then same-box A/B of the exchange arrangement and the CG arrangement
mkdir -p gpurun_out
run() {  # name, env..., then bench args
  name=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 8 --no-cpu --steps 500 --warmup 20 --e2e-steps 5 $EXTRA > gpurun_out/r2f_$name.json 2> gpurun_out/r2f_$name.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2f_$name.json').read().strip().splitlines()[-1])
    r=d['roofline']; p=d['config'].get('parity') or {}
    print('$name', 'explicit ms %.5f'%d['ms_per_step'], 'tile', r.get('tile_ms'), 'bnd', r.get('boundary_ms'), 'e2e %.3f'%d['e2e']['ms_per_step'], {k:round(v['ms_per_iter'],5) for k,v in (d.get('solvers') or {}).items()}, 'parity', p.get('status'), {k:v for k,v in p.items() if k.startswith('rel_')})
except Exception as e:
    print('$name FAILED', e); print(open('gpurun_out/r2f_$name.err').read()[-1500:])
PY
}
EXTRA="" run default FEMB_X=1
EXTRA="--no-parity" run late FEMB_PEER_MODE=late
EXTRA="--no-parity" run cg0 FEMB_CG_SINGLE_REDUCTION=0
EXTRA="--no-parity --workload c4_explicit" run late_pe FEMB_PEER_MODE=late FEMB200_LIB=$PWD/variants/persist_explicit.so
