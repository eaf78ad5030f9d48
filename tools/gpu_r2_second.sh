#!/bin/bash
# Round 2, second visit (2 GPUs): the whole GPU suite incl. the 2-rank tests, the new default bench line at N=1 and N=2,
# and the reference arm.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2b_pytest.log; tail -5 gpurun_out/r2b_pytest.log
timeout 900 python bench.py --steps 200 --warmup 10 > gpurun_out/r2b_bench1.json 2> gpurun_out/r2b_bench1.err; echo "bench1 rc=$?"; tail -c 1500 gpurun_out/r2b_bench1.json; tail -5 gpurun_out/r2b_bench1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 200 --warmup 10 > gpurun_out/r2b_bench2.json 2> gpurun_out/r2b_bench2.err; echo "bench2 rc=$?"; tail -c 1500 gpurun_out/r2b_bench2.json; tail -5 gpurun_out/r2b_bench2.err
timeout 900 python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r2b_ref.json 2> gpurun_out/r2b_ref.err; echo "ref rc=$?"; tail -c 800 gpurun_out/r2b_ref.json
