#!/bin/bash
# development tool: 2-GPU run — loop-back/multi-rank tests, weak-scaling bench with the interleaved boundary tiles and with the old order
o=gpurun_out; mkdir -p $o
(timeout 900 python -m pytest tests/test_gpu_loopback.py tests/test_gpu_multi.py tests/test_gpu_plugin.py tests/test_gpu_ksp.py -m gpu -q) > $o/r02_gputest_2gpu_c.log 2>&1; echo "pytest rc=$?"; tail -4 $o/r02_gputest_2gpu_c.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
$T 29621 bench.py --gpus 2 > $o/r02_bench_n2_order1.json 2> $o/r02_bench_n2_order1.err; echo "bench order1 rc=$?"
B200_MPI_FUSED_ORDER=0 $T 29622 bench.py --gpus 2 --no-cpu-baseline --no-plugin --no-solve --strong-nx 0 > $o/r02_bench_n2_order0.json 2> $o/r02_bench_n2_order0.err; echo "bench order0 rc=$?"
$T 29623 bench.py --gpus 2 --no-cpu-baseline --no-plugin --no-solve --strong-nx 0 > $o/r02_bench_n2_order1b.json 2> $o/r02_bench_n2_order1b.err; echo "bench order1b rc=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob("$o/r02_bench_n2_order*.json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        print(f, "ms/step %.4f per-launch min %.4f med %.4f cg %.4f ms"%(d["ms_per_step"],d["per_launch"]["min_ms"],d["per_launch"]["median_ms"],d["kspcg"]["fused"]["ms_per_iter"]), d.get("plugin_on_ranks",{}).get("cg",{}).get("y_norm1_rel_diff"))
    except Exception as e: print(f,"unreadable",e)
PY
