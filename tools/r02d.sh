o=gpurun_out
python -m pytest tests/test_gpu_loopback.py -q > $o/r02d_loopback.log 2>&1; echo "loopback rc=$?"; tail -2 $o/r02d_loopback.log
python -m pytest tests/test_gpu_plugin_mpi.py -q > $o/r02d_plugin_mpi.log 2>&1; echo "plugin_mpi rc=$?"; tail -2 $o/r02d_plugin_mpi.log
python -m pytest tests/test_gpu_plugin.py -q -k zerorows > $o/r02d_plugin_zr.log 2>&1; echo "plugin zr rc=$?"
T="timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port"
$T 29611 bench.py --gpus 2 --steps 50 --warmup 5 > $o/r02d_bench_n2.json 2> $o/r02d_bench_n2.err; echo "bench2 rc=$?"
python tools/configs_bench.py c1 > $o/r02d_c1_persist.jsonl 2> $o/r02d_c1_persist.err; echo "c1 rc=$?"
B200_CG_PERSISTENT=0 python tools/configs_bench.py c1 > $o/r02d_c1_nopersist.jsonl 2>> $o/r02d_c1_persist.err; echo "c1 np rc=$?"
cat $o/r02d_c1_persist.jsonl $o/r02d_c1_nopersist.jsonl | cut -c1-400
