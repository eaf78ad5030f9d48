#!/bin/bash
# What round 2 could not run (its GPU-minutes were spent before this code was written) -- one command for whoever has
# GPU time next:   gpurun --timeout 1500 -- 'bash tools/gpu_r2_deferred.sh'
#   1. GPU parity suite on the NEW tile tables (default) -- must stay green
#   2. the reference-order sweep's launch plumbing as plain tests (FEMB_SWEEP_STRICT=1 removes their xfail mark)
#   3. same-box A/B: round-1 tables vs each step of the bank-conflict work, three workloads
#   4. ncu of the explicit tile kernel on the new tables: the shared-memory wavefront / bank-conflict counters that
#      tools/bank_model.py predicts (35.31 M / 10.99 M on the round-1 tables -> modelled -7.5 M)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log; tail -4 gpurun_out/pytest_gpu.log
FEMB_SWEEP_STRICT=1 timeout 600 python -m pytest tests/test_zz_gpu_sweep.py -m gpu -q > gpurun_out/pytest_sweep.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_sweep.log; tail -6 gpurun_out/pytest_sweep.log
ab() {   # name, env...
  local name=$1; shift
  for w in c4_explicit c4_cg c4_jacobi; do
    S=200; [[ $w == *explicit ]] && S=500
    env "$@" timeout 300 python bench.py --no-cpu --workload $w --steps $S --warmup 10 --e2e-steps 3 2> gpurun_out/ab.err | tail -1 | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('$name $w', 'ms/step %.4f'%d['ms_per_step'], 'frac %.3f'%r['frac'], 'tile_ms', r.get('tile_ms'), 'bnd_ms', r.get('boundary_ms'), 'value %.4g'%d['value'])
except Exception as e: print('$name $w FAILED', e)
" | tee -a gpurun_out/ab_tables.log
  done
}
ab round1_tables   FEMB_LIST_ORDER=greedy FEMB_SLOT_ORDER=id FEMB_VERTEX_ORDER=id
ab lists_matching  FEMB_SLOT_ORDER=id FEMB_VERTEX_ORDER=id
ab lists+slots     FEMB_VERTEX_ORDER=id
ab default         FEMB_X=1
ab round1_tables_again FEMB_LIST_ORDER=greedy FEMB_SLOT_ORDER=id FEMB_VERTEX_ORDER=id
M=l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum,l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,gpu__time_duration.sum,smsp__inst_executed.sum
for t in "FEMB_X=1" "FEMB_LIST_ORDER=greedy FEMB_SLOT_ORDER=id FEMB_VERTEX_ORDER=id"; do
  tag=$(echo "$t" | tr ' =' '__')
  env $t timeout 600 ncu --metrics $M --clock-control none -k regex:k_tile -c 6 --csv --log-file gpurun_out/ncu_smem_$tag.csv \
    python bench.py --no-cpu --workload c4_explicit --steps 2 --warmup 1 --e2e-steps 0 > gpurun_out/ncu_smem_$tag.log 2>&1
done
