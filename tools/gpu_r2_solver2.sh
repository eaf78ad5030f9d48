#!/bin/bash
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2e_pytest.log; tail -15 gpurun_out/r2e_pytest.log
timeout 600 python bench.py --no-cpu --steps 200 --warmup 10 > gpurun_out/r2e_bench1.json 2> gpurun_out/r2e_bench1.err; echo "bench rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2e_bench1.json').read().strip().splitlines()[-1]); print('N=1', d['ms_per_step'], {k:(v['ms_per_iter'], v['roofline']['iteration_frac'], v['e2e']['value'], v['e2e']['iterations_per_solve']) for k,v in d['solvers'].items()})"
if [ "$NG" -ge 2 ]; then
for mode in fused late; do
FEMB_PEER_MODE=$mode timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 500 --warmup 20 --no-cpu > gpurun_out/r2e_bench2_$mode.json 2> gpurun_out/r2e_bench2_$mode.err; echo "bench2 $mode rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2e_bench2_$mode.json').read().strip().splitlines()[-1]); print('N=2 $mode', d['ms_per_step'], d['config']['parity']['status'], {k:v['ms_per_iter'] for k,v in d['solvers'].items()})"
done
fi
