#!/bin/bash
# N-GPU visit: the 2-rank slab check (both transports), benches at N GPUs (every command under its own timeout)
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -12 gpurun_out/pytest_gpu2.log | cut -c1-600
for comm in peer nccl; do for w in c4_explicit c4_cg c4_jacobi; do
  S=200; [ $w = c4_explicit ] && S=1000
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus $N --workload $w --steps $S --warmup 10 --no-cpu --comm $comm > gpurun_out/b${N}_${comm}_$w.json 2> gpurun_out/b${N}_${comm}_$w.err
  tail -1 gpurun_out/b${N}_${comm}_$w.json | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('N=$N $comm $w ms/step %.4f value %.4g comm=%s launches %d' % (d['ms_per_step'], d['value'], d['config']['comm'], d['gpu_launches']))
except Exception as e: print('N=$N $comm $w FAILED', e)
"; grep -i "error\|unavailable\|Traceback" gpurun_out/b${N}_${comm}_$w.err | head -3
done; done
