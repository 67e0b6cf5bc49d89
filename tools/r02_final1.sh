#!/bin/bash
# the closing single-GPU run of round 2: what the driver runs (GPU tests, smoke, both bench arms) + the launch list
o=gpurun_out; mkdir -p $o
(time timeout 1500 python -m pytest tests -m gpu -q) > $o/r02_final_gputest.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02_final_gputest.log
python __graft_entry__.py smoke > $o/r02_final_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 $o/r02_final_smoke.log
(time python bench.py) > $o/r02_final_bench_n1.json 2> $o/r02_final_bench_n1.err; echo "bench rc=$?"
(time python bench.py --impl reference) > $o/r02_final_bench_ref_n1.json 2> $o/r02_final_bench_ref_n1.err; echo "bench ref rc=$?"; cut -c1-600 $o/r02_final_bench_ref_n1.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $o/r02_final_bench_launches.csv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-solve --strong-nx 0 > $o/r02_final_bench_ncu.log 2>&1; echo "launch list rc=$?"
python - <<PY
import json
d=json.loads([l for l in open("$o/r02_final_bench_n1.json") if l.startswith("{")][-1])
print("value %.0f ms/step %.4f frac %.3f cg fused %.4f op %.4f strong %s e2e %.0f launches %s parity %s cpu %s"%(d["value"],d["ms_per_step"],d["roofline"]["frac"],d["kspcg"]["fused"]["ms_per_iter"],d["kspcg"]["op_by_op"]["ms_per_iter"],d["kspcg_strong"]["iter_per_s"],d["e2e"]["value"],d["gpu_launches"],d["parity"]["ok"],d["cpu_baseline"]["value"]))
PY
