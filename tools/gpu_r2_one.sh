#!/bin/bash
# "one tile kernel per pass" exchange: 2-rank parity tests, then (N ranks) A/B against the split form + timeline
mkdir -p gpurun_out
N=${1:-2}
[ -n "$SKIP_TESTS" ] || { timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2i_pytest.log; tail -4 gpurun_out/r2i_pytest.log; }
run() {
  name=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29741 bench.py --gpus $N --no-cpu --steps 500 --warmup 20 --e2e-steps 5 $EXTRA > gpurun_out/r2i_n${N}_$name.json 2> gpurun_out/r2i_n${N}_$name.err
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/r2i_n${N}_$name.json').read().strip().splitlines()[-1])
    p=d['config'].get('parity') or {}
    print('N=$N $name', 'explicit ms %.5f'%d['ms_per_step'], 'e2e %.3f'%d['e2e']['ms_per_step'], {k:round(v['ms_per_iter'],5) for k,v in (d.get('solvers') or {}).items()}, 'parity', p.get('status'), {k:v for k,v in p.items() if k.startswith('rel_')})
except Exception as e:
    print('$name FAILED', e); print(open('gpurun_out/r2i_n${N}_$name.err').read()[-1500:])
PY
}
EXTRA="" run split FEMB_X=1
EXTRA="--no-parity" run one FEMB_PEER_MODE=one FEMB200_LIB=$PWD/finite-element-method_b200/femb200/libfemb200_diag.so
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29743 tools/trace_steps.py --out gpurun_out/trace_split_n$N.txt > gpurun_out/trace_split_n$N.log 2>&1; echo "trace rc=$?"
FEMB_PEER_MODE=one timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29745 tools/trace_steps.py --out gpurun_out/trace_one_n$N.txt > gpurun_out/trace_one_n$N.log 2>&1; echo "trace one rc=$?"
