#!/bin/bash
# development tool: reduction grid — one wave of resident CTAs against the default cap (two waves)
for w in 0 1 2; do echo "B200_RED_WAVES=$w"; B200_RED_WAVES=$w python tools/vec_ncu_driver.py --time 2>/dev/null | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l)
        if d['op'] in ('VecDot','VecNorm2','VecMDot8'): print('  %-10s %.4f ms %.0f GB/s (%.3f)'%(d['op'],d['ms'],d['gbs'],d['frac_of_measured_peak']))"; done
