o=gpurun_out
python -m pytest tests/test_gpu_loopback.py -x -q > $o/r02c_loopback.log 2>&1; echo "loopback rc=$?"
python -m pytest tests/test_gpu_multi.py -x -q > $o/r02c_multi.log 2>&1; echo "multi rc=$?"; tail -3 $o/r02c_multi.log
T="timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
$T 29611 bench.py --gpus 2 --steps 50 --warmup 5 > $o/r02c_bench_n2.json 2> $o/r02c_bench_n2.err; echo "bench2 rc=$?"
B200_MPI_FUSED_OCC8=0 $T 29612 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline --no-plugin --strong-nx 0 > $o/r02c_bench_n2_occ6.json 2> $o/r02c_bench_n2_occ6.err; echo "bench2 occ6 rc=$?"
B200_MPI_FUSED=0 $T 29613 bench.py --gpus 2 --steps 50 --warmup 5 --no-cpu-baseline --no-plugin --strong-nx 0 > $o/r02c_bench_n2_unfused.json 2> $o/r02c_bench_n2_unfused.err; echo "bench2 unfused rc=$?"
$T 29614 bench.py --impl reference --gpus 2 --steps 10 --warmup 3 > $o/r02c_bench_ref_n2.json 2> $o/r02c_bench_ref_n2.err; echo "ref2 rc=$?"
