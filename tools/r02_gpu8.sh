#!/bin/bash
# development tool: the 8-GPU confirmation run of round 2 (parity on 8 real ranks, plugin under MPI on 3/4/8 ranks, weak + strong bench)
o=gpurun_out; mkdir -p $o
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29590 tests/multi_gpu_worker.py p2p > $o/r02_multi8_p2p.log 2>&1; echo "parity8 rc=$?"; grep -o "MULTI_GPU_RESULT.*" $o/r02_multi8_p2p.log | cut -c1-600
(timeout 700 python -m pytest tests/test_gpu_plugin_mpi.py -m gpu -q -k "3 or 4 or 8") > $o/r02_gputest_8gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $o/r02_gputest_8gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29608 bench.py --gpus 8 > $o/r02_bench_n8.json 2> $o/r02_bench_n8.err; echo "bench8 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29604 bench.py --gpus 4 --no-plugin > $o/r02_bench_n4.json 2> $o/r02_bench_n4.err; echo "bench4 rc=$?"
python - <<PY
import json,glob
for f in sorted(glob.glob("$o/r02_bench_n[48].json")):
    try:
        d=json.loads([l for l in open(f) if l.startswith("{")][-1])
        s=d.get("kspcg_strong",{})
        print(f, "value %.0f ms/step %.4f med %.4f frac %.3f cg fused %.4f ms strong %s it/s e2e %.0f parity %s"%(d["value"],d["ms_per_step"],d["per_launch"]["median_ms"],d["roofline"]["frac"],d["kspcg"]["fused"]["ms_per_iter"],s.get("iter_per_s"),d["e2e"]["value"],d["parity"]), json.dumps(d.get("plugin_on_ranks"))[:500])
    except Exception as e: print(f, "unreadable", e)
PY
tail -3 $o/r02_bench_n8.err
