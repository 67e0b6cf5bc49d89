#!/bin/bash
# development tool: the single-GPU evidence run (tests, bench, reference arm, ncu launch list + full captures)
o=gpurun_out; tag=${1:-r01}
(time python -m pytest tests -m gpu -q) > $o/${tag}_pytest.log 2>&1; tail -3 $o/${tag}_pytest.log
python bench.py 2> $o/${tag}_bench.err | grep "^{" > $o/${tag}_bench_n1.json; echo bench rc=$?; cut -c1-600 $o/${tag}_bench_n1.json
(time python bench.py --impl reference --steps 5 --warmup 1) 2> $o/${tag}_ref.err | grep "^{" > $o/${tag}_bench_reference.json; cut -c1-400 $o/${tag}_bench_reference.json
# launch list of the same bench command (short)
python bench.py --steps 5 --warmup 3 --cg-iters 10 --no-cpu-baseline > $o/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $o/${tag}_bench_launches.csv python bench.py --steps 5 --warmup 3 --cg-iters 10 --no-cpu-baseline > $o/${tag}_ncu1.log 2>&1
# full captures: the MatMult kernel and the fused CG kernels
python tools/ncu_driver.py lap7 256 cg > $o/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:spmv_tile -s 2 -c 1 -f -o $o/${tag}_spmv_lap7 python tools/ncu_driver.py lap7 256 > $o/${tag}_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:cg_update -s 3 -c 1 -f -o $o/${tag}_cg_update python tools/ncu_driver.py lap7 256 cg > $o/${tag}_ncu3.log 2>&1
tail -2 $o/${tag}_ncu2.log $o/${tag}_ncu3.log
