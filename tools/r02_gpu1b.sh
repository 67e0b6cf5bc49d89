#!/bin/bash
# development tool: single-GPU A/B runs of round 2 (persistent CG variants on configs[0]; gather load mode on the power-law sample)
o=gpurun_out; mkdir -p $o
run_c1() { tag=$1; shift; env "$@" python tools/configs_bench.py c1 2>> $o/r02b_c1.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$tag', 'cg fused %.2f us/iter (%.0f it/s)  op-by-op %.2f us  matmult %.2f us  upload %.3f s'%(1e3*d['kspcg_fused']['ms_per_iter'], d['kspcg_fused']['iter_per_s'], 1e3*d['kspcg_op_by_op']['ms_per_iter'], 1e3*d['matmult_ms'], d['upload_s']))
"; }
run_c1 persist_default      B200_X=0
run_c1 persist_keepl2       B200_CG_PERSIST_KEEPL2=1
run_c1 persist_occ6         B200_CG_PERSIST_OCC=6
run_c1 persist_occ2         B200_CG_PERSIST_OCC=2
run_c1 persist_keepl2_occ6  B200_CG_PERSIST_KEEPL2=1 B200_CG_PERSIST_OCC=6
run_c1 nopersist            B200_CG_PERSISTENT=0
python tools/plugin_e2e.py 1000 > $o/r02b_plugin_e2e_ex2_1000.json 2>> $o/r02b_c1.err; python -c "
import json; d=json.load(open('$o/r02b_plugin_e2e_ex2_1000.json'))
for k in ('reference','b200','b200_fused'): print('plugin_e2e',k,d[k].get('iterations'),d[k].get('kspsolve_s'),d[k].get('iter_per_s'))"
B200_SPMV_GATHER_CG=0 python tools/spmv_bench.py plaw 2>&1 | grep "^plaw" | sed 's/^/gather_nc /'
B200_SPMV_GATHER_CG=1 python tools/spmv_bench.py plaw 2>&1 | grep "^plaw" | sed 's/^/gather_cg /'
python tools/configs_bench.py c2 2>> $o/r02b_c1.err | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('C2 upload %.3f s matmult %.4f ms cg %.4f ms'%(d['upload_s'], d['matmult_ms'], d['kspcg_fused']['ms_per_iter']))"
for c in 8 7 6 5; do B200_SPMV_CTAS=$c python tools/spmv_bench.py lap7 2>&1 | grep "^lap7" | sed "s/^/ctas_per_sm=$c /"; done
