#!/bin/bash
# development tool: 2-GPU check of a fused-kernel change — loop-back + multi-rank tests, short weak-scaling bench
o=gpurun_out; mkdir -p $o
(timeout 600 python -m pytest tests/test_gpu_loopback.py tests/test_gpu_multi.py -m gpu -q) > $o/r02_gputest_2gpu_e.log 2>&1; echo "pytest rc=$?"; tail -3 $o/r02_gputest_2gpu_e.log
T="timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
for k in 1 2; do
$T 2964$k bench.py --gpus 2 --no-cpu-baseline --no-plugin --no-solve --strong-nx 0 > $o/r02_bench_n2_v4_$k.json 2> $o/r02_bench_n2_v4_$k.err; echo "bench rc=$?"
done
python - <<PY
import json,glob
for f in sorted(glob.glob("$o/r02_bench_n2_v4_*.json")):
    d=json.loads([l for l in open(f) if l.startswith("{")][-1])
    print(f, "ms/step %.4f per-launch min %.4f med %.4f cg %.4f ms parity %s"%(d["ms_per_step"],d["per_launch"]["min_ms"],d["per_launch"]["median_ms"],d["kspcg"]["fused"]["ms_per_iter"],d["parity"]["ok"]))
PY
