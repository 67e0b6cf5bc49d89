#!/bin/bash
# development tool: single-GPU check of the three-launch CG iteration (product's last CTA takes the scalar step)
o=gpurun_out; mkdir -p $o
(timeout 900 python -m pytest tests/test_gpu_ksp.py tests/test_gpu_mat.py tests/test_gpu_plugin.py tests/test_gpu_fullsize.py -m gpu -q) > $o/r02_gputest_1gpu_f.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02_gputest_1gpu_f.log
python tools/c1_sweep.py > $o/r02_c1_sweep_v3.txt 2>&1; cat $o/r02_c1_sweep_v3.txt
python tools/configs_bench.py c2 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('C2 upload %.3f s matmult %.4f ms cg fused %.4f ms op %.4f ms'%(d['upload_s'], d['matmult_ms'], d['kspcg_fused']['ms_per_iter'], d['kspcg_op_by_op']['ms_per_iter']))"
python tools/plugin_e2e.py 1000 > $o/r02_plugin_e2e_v3.json 2>/dev/null; python -c "
import json; d=json.load(open('$o/r02_plugin_e2e_v3.json'))
for k in ('b200','b200_fused'): print('plugin_e2e',k,d[k].get('iterations'),d[k].get('kspsolve_s'),d[k].get('iter_per_s'),d[k].get('norm_of_error'))"
