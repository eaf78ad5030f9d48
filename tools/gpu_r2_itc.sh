#!/bin/bash
# ITC (boundary vertices finished inside the tile kernel): full GPU suite on the new default, then same-box A/B
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/r2c_pytest.log; tail -5 gpurun_out/r2c_pytest.log
FEMB200_LIB=$PWD/variants/persist_explicit.so timeout 900 python -m pytest tests/test_gpu_explicit.py tests/test_gpu_ref_golden.py tests/test_gpu_fullsize.py -x -q > gpurun_out/r2c_pytest_pe.log 2>&1; echo "rc=$?" >> gpurun_out/r2c_pytest_pe.log; tail -3 gpurun_out/r2c_pytest_pe.log
for rep in 1 2; do for so in base persist_explicit; do for itc in 0 1; do for w in c4_explicit c3_explicit; do
  S=500; [[ $w == c3* ]] && S=2000
  FEMB_ITC=$itc FEMB200_LIB=$PWD/variants/$so.so timeout 300 python bench.py --no-cpu --workload $w --steps $S --warmup 20 --e2e-steps 3 2>> gpurun_out/r2c_ab.err | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); r=d['roofline']
print('$so itc=$itc $w rep$rep', 'ms/step %.5f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'step_frac %.3f'%r['step_frac'], 'tile_ms', r.get('tile_ms'), 'bnd_ms', r.get('boundary_ms'))
" | tee -a gpurun_out/r2c_ab.log
done; done; done; done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus 2 --steps 500 --warmup 20 --no-cpu > gpurun_out/r2c_bench2.json 2> gpurun_out/r2c_bench2.err; echo "bench2 rc=$?"
python -c "
import json; d=json.loads(open('gpurun_out/r2c_bench2.json').read().strip().splitlines()[-1]); print('N=2', d['ms_per_step'], d['config']['parity'], {k:v['ms_per_iter'] for k,v in d['solvers'].items()})"
