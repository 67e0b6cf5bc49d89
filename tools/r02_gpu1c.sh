#!/bin/bash
# development tool: single-GPU run after the register diet — full GPU test suite, small-problem sweep, kernels, assembly
o=gpurun_out; mkdir -p $o
(time timeout 1200 python -m pytest tests -m gpu -q) > $o/r02_gputest_1gpu_c.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02_gputest_1gpu_c.log
python tools/c1_sweep.py cfg > $o/r02_c1_sweep.txt 2>&1; echo "c1 sweep rc=$?"; cat $o/r02_c1_sweep.txt
python tools/spmv_bench.py lap7 s27 bs3 plaw > $o/r02_spmv_bench.txt 2>&1; cat $o/r02_spmv_bench.txt
python tools/configs_bench.py c2 c5 > $o/r02_configs_c2c5.jsonl 2>$o/r02_configs_c2c5.err; python - <<PY
import json
for l in open("$o/r02_configs_c2c5.jsonl"):
    if l.startswith("{"):
        d=json.loads(l); print(d["config"], "upload %.3f s matmult %.4f ms frac %.3f cg fused %.4f ms/iter op-by-op %.4f"%(d["upload_s"], d["matmult_ms"], d["matmult_frac_of_measured_peak"], d["kspcg_fused"]["ms_per_iter"], d["kspcg_op_by_op"]["ms_per_iter"]))
PY
python tools/assembly_bench.py > $o/r02_assembly_hex128.json 2> $o/r02_assembly.err; echo "assembly rc=$?"; cut -c1-1200 $o/r02_assembly_hex128.json
