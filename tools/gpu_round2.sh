#!/bin/bash
# 2-GPU visit: the 2-rank slab check, implicit + explicit benches at N=2 (every command under its own timeout)
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu2.log
tail -8 gpurun_out/pytest_gpu2.log
for w in c4_explicit c4_cg c4_jacobi; do
  S=200; [ $w = c4_explicit ] && S=1000
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 2 --workload $w --steps $S --warmup 10 --no-cpu > gpurun_out/b2_$w.json 2> gpurun_out/b2_$w.err; tail -c 700 gpurun_out/b2_$w.json; tail -3 gpurun_out/b2_$w.err
done
