#!/bin/bash
# scaling visit on an N-GPU box: bench at the given rank counts (peer-memory transport), every run under a timeout
mkdir -p gpurun_out
for N in "$@"; do for w in c4_explicit c4_cg c4_jacobi; do
  S=200; [ $w = c4_explicit ] && S=1000
  if [ $N = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621"; fi
  timeout 240 $L bench.py --gpus $N --workload $w --steps $S --warmup 10 --no-cpu --e2e-steps 3 > gpurun_out/s${N}_$w.json 2> gpurun_out/s${N}_$w.err
  tail -1 gpurun_out/s${N}_$w.json | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('N=$N $w ms/step %.4f value %.4g comm=%s tiles %d launches %d' % (d['ms_per_step'], d['value'], str(d['config']['comm'])[:12], d['config']['n_tiles_rank0'], d['gpu_launches']))
except Exception as e: print('N=$N $w FAILED', e)
" | tee -a gpurun_out/scale.log; grep -i "error\|unavailable\|Traceback" gpurun_out/s${N}_$w.err | head -3
done; done
