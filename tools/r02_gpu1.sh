#!/bin/bash
# development tool: the single-GPU evidence run of round 2 (outputs under gpurun_out/, summaries copied to profiles/ by hand)
o=gpurun_out; mkdir -p $o
python tools/configs_bench.py c1 c2 > $o/r02_configs_c1c2.jsonl 2> $o/r02_configs_c1c2.err; echo "c1c2 rc=$?"; cut -c1-600 $o/r02_configs_c1c2.jsonl
B200_CG_PERSISTENT=0 python tools/configs_bench.py c1 > $o/r02_configs_c1_nopersist.jsonl 2>> $o/r02_configs_c1c2.err; echo "c1 nopersist rc=$?"
python tools/plugin_e2e.py 1000 > $o/r02_plugin_e2e_ex2_1000.json 2> $o/r02_plugin_e2e.err; echo "plugin_e2e rc=$?"; cut -c1-900 $o/r02_plugin_e2e_ex2_1000.json
# launch list of the default bench command (after it ran clean in the previous call)
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/r02_bench_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-solve --strong-nx 0 > $o/r02_bench_ncu.log 2>&1; echo "launch list rc=$?"
# C4 at its full size: timing, then ncu --set full of one product of the final kernel (PARK + chunks + strips)
export B200_CACHE_DIR=/tmp
python tools/configs_bench.py c4 > $o/r02_configs_c4.jsonl 2> $o/r02_configs_c4.err; echo "c4 rc=$?"; cut -c1-700 $o/r02_configs_c4.jsonl
python tools/ncu_driver.py plaw 10000000 > $o/r02_plaw_plain.log 2>&1; echo "plaw plain rc=$?"; tail -2 $o/r02_plaw_plain.log
ncu --set full --clock-control none --import-source on -k regex:spmv_tile --launch-skip 8 -c 4 -o $o/r02_spmv_plaw_10M python tools/ncu_driver.py plaw 10000000 > $o/r02_plaw_ncu.log 2>&1; echo "plaw ncu rc=$?"
ls -la $o | tail -20
