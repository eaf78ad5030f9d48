#!/bin/bash
mkdir -p gpurun_out
for dbg in 0 1 2 3; do
  FEMB200_DEBUG_COMM=$dbg timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 2 --workload c4_explicit --steps 1000 --warmup 10 --no-cpu --e2e-steps 3 2> gpurun_out/dbg.err | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('dbg=$dbg ms/step %.4f' % d['ms_per_step'], d['roofline']['tile_ms'], d['roofline']['boundary_ms'])"
done
