#!/bin/bash
# development tool: the multi-GPU evidence run on an 8-GPU box (parity at 8 ranks, weak-scaling bench, strong-scaling CG)
o=gpurun_out; tag=${1:-r01}
run() { # n transport extra...
  n=$1; t=$2; shift 2
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 50 --warmup 5 --transport $t "$@"
}
for t in ${PARITY_TRANSPORTS:-p2p}; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29590 tests/multi_gpu_worker.py $t > $o/${tag}_multi8_$t.log 2>&1; echo "parity8 $t rc=$?"; grep -o "MULTI_GPU_RESULT.*" $o/${tag}_multi8_$t.log | cut -c1-300
done
for n in 8 4 2; do
  run $n p2p --strong-nx 512 2> $o/${tag}_bench_n${n}_p2p.err | grep "^{" > $o/${tag}_bench_n${n}_p2p.json; echo "bench $n p2p rc=$?"
done
run 8 nccl 2> $o/${tag}_bench_n8_nccl.err | grep "^{" > $o/${tag}_bench_n8_nccl.json; echo "bench 8 nccl rc=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob("$o/${tag}_bench_n*_*.json")):
    try:
        d=json.load(open(f))
        s=d.get("kspcg_strong",{})
        print(f, "ms/step %.4f value %.0f frac %.3f cg %.4f ms strong %s it/s mm %s ms"%(d["ms_per_step"],d["value"],d["roofline"]["frac"],d["kspcg"]["fused"]["ms_per_iter"],s.get("iter_per_s"),s.get("matmult_ms")))
    except Exception as e: print(f, "unreadable", e)
PY
