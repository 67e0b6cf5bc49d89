#!/bin/bash
# development tool: sweep the tile-kernel launch shapes / lanes-per-row on the BASELINE matrix shapes
out=gpurun_out/sweep.log; : > $out
for m in lap7 s27 bs3 lap2d; do echo "== $m default" >> $out; timeout 120 python tools/spmv_bench.py $m >> $out 2>&1; done
for pk in 0 1; do for g in 1 2 4 8; do for l in 128 256 512; do echo "== plaw park $pk G $g long $l" >> $out; B200_SPMV_PARK=$pk B200_SPMV_G=$g B200_SPMV_LONG=$l timeout 120 python tools/spmv_bench.py plaw >> $out 2>&1; done; done; done
for c in 1 3 4; do echo "== plaw default cfg $c" >> $out; B200_SPMV_CFG=$c timeout 120 python tools/spmv_bench.py plaw >> $out 2>&1; done
