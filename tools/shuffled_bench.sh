#!/bin/bash
# development aid: C2 dense / shuffled cull time for tuning variants
for lib in ${1:-product}; do
  if [ "$lib" = product ]; then unset RVE_B200_LIB; else export RVE_B200_LIB=$PWD/$lib; fi
  python tools/stage_bench.py 2>/dev/null | head -3 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$lib', d['workload'][:60], round(d['ms_median'],4))"
done
