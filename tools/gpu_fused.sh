#!/bin/bash
# 2-GPU visit for the fused halo exchange: slab parity (both transports), then per-rank-size proxies:
# x-cells 36 on 2 ranks = the slab a rank owns at N=8 on C4; FEMB_PEER_FUSED=0 is the old send/recv-kernel path.
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -6 gpurun_out/pytest_gpu2.log | cut -c1-400
L="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621"
run() {  # fused x_cells workload steps
  FEMB_PEER_FUSED=$1 timeout 200 $L bench.py --gpus 2 --workload $3 --steps $4 --warmup 10 --no-cpu --e2e-steps 3 --x-cells $2 > gpurun_out/f$1_x$2_$3.json 2> gpurun_out/f$1_x$2_$3.err
  tail -1 gpurun_out/f$1_x$2_$3.json | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('fused=$1 x=$2 $3 ms/step %.4f launches %d' % (d['ms_per_step'], d['gpu_launches']))
except Exception as e: print('fused=$1 x=$2 $3 FAILED', e)
" | tee -a gpurun_out/fused.log
}
run 1 0 c4_cg 200; run 1 0 c4_jacobi 200
run 1 72 c4_cg 200
export FEMB_PEER_FUSED_WAVES=16
run 1 72 c4_cg 200
