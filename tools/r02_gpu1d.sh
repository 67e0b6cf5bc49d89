#!/bin/bash
# development tool: single-GPU run — CG graph replay, plugin end to end on configs[0], ncu captures of the headline kernels
o=gpurun_out; mkdir -p $o
(timeout 600 python -m pytest tests/test_gpu_ksp.py tests/test_gpu_plugin.py -m gpu -q) > $o/r02_gputest_1gpu_d.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02_gputest_1gpu_d.log
python tools/c1_sweep.py > $o/r02_c1_sweep_graph.txt 2>&1; cat $o/r02_c1_sweep_graph.txt
python tools/plugin_e2e.py 1000 > $o/r02_plugin_e2e_ex2_1000.json 2> $o/r02_plugin_e2e.err; python -c "
import json; d=json.load(open('$o/r02_plugin_e2e_ex2_1000.json'))
for k in ('reference','b200','b200_fused'): print('plugin_e2e',k,d[k].get('iterations'),d[k].get('kspsolve_s'),d[k].get('iter_per_s'),d[k].get('norm_of_error'))"
python tools/configs_bench.py c1 c3 > $o/r02_configs_c1c3.jsonl 2> $o/r02_configs_c1c3.err; cut -c1-250 $o/r02_configs_c1c3.jsonl
# ncu --set full of the headline kernels (after the same commands ran clean above / in earlier calls)
ncu --set full --clock-control none --import-source on -k regex:spmv_tile --launch-skip 2 -c 1 -o $o/r02_spmv_lap7_256 python tools/ncu_driver.py lap7 256 > $o/r02_ncu_lap7.log 2>&1; echo "ncu lap7 rc=$?"
ncu --set full --clock-control none -k regex:cg_update --launch-skip 2 -c 1 -o $o/r02_cg_update_256 python tools/ncu_driver.py lap7 256 cg > $o/r02_ncu_cg.log 2>&1; echo "ncu cg rc=$?"
ls -la $o/*.ncu-rep
