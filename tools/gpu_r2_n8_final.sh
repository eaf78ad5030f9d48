#!/bin/bash
# final N-rank record of round 2: the default bench line (with the N>1 parity leg) and the kernel timeline
mkdir -p gpurun_out
N=${1:-8}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29751 bench.py --gpus $N --no-cpu --steps 500 --warmup 20 > gpurun_out/r2x_bench_n$N.json 2> gpurun_out/r2x_bench_n$N.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r2x_bench_n$N.json').read().strip().splitlines()[-1])
p=d['config'].get('parity') or {}
print('N=$N', 'explicit ms %.5f'%d['ms_per_step'], 'e2e %.3f'%d['e2e']['ms_per_step'], {k:round(v['ms_per_iter'],5) for k,v in (d.get('solvers') or {}).items()}, 'parity', p.get('status'))
PY
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29753 tools/trace_steps.py --out gpurun_out/r2x_trace_n$N.txt > gpurun_out/r2x_trace_n$N.log 2>&1; echo "trace rc=$?"
