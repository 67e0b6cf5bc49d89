#!/bin/bash
# development tool: 2-GPU run after the kernel restructuring — tests that drive the fused MPIAIJ kernel, weak-scaling bench, small-problem CG
o=gpurun_out; mkdir -p $o
(timeout 1200 python -m pytest tests/test_gpu_loopback.py tests/test_gpu_multi.py tests/test_gpu_ksp.py tests/test_gpu_mpiaij.py tests/test_gpu_assembly.py tests/test_gpu_plugin_mpi.py -m gpu -q) > $o/r02_gputest_2gpu_d.log 2>&1; echo "pytest rc=$?"; tail -6 $o/r02_gputest_2gpu_d.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
$T 29631 bench.py --gpus 2 > $o/r02_bench_n2_v3.json 2> $o/r02_bench_n2_v3.err; echo "bench n2 rc=$?"
python bench.py --no-cpu-baseline > $o/r02_bench_n1_v3.json 2> $o/r02_bench_n1_v3.err; echo "bench n1 rc=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob("$o/r02_bench_n?_v3.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "value %.0f ms/step %.4f per-launch min %.4f med %.4f cg fused %.4f op %.4f strong %s e2e %.0f parity %s"%(d["value"],d["ms_per_step"],d["per_launch"]["min_ms"],d["per_launch"]["median_ms"],d["kspcg"]["fused"]["ms_per_iter"],d["kspcg"]["op_by_op"]["ms_per_iter"],d.get("kspcg_strong",{}).get("iter_per_s"),d["e2e"]["value"],d["parity"]["ok"]), d.get("plugin_on_ranks",{}).get("cg",{}).get("y_norm1_rel_diff"))
    except Exception as e: print(f,"unreadable",e)
PY
python tools/c1_sweep.py > $o/r02_c1_sweep_graph.txt 2>&1; cat $o/r02_c1_sweep_graph.txt
