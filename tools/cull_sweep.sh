#!/bin/bash
# development aid: kernel_ms of the C2 cull kernel for tuning variants (build/*.so) x RVE_CULL_DEBUG settings
# usage: tools/cull_sweep.sh "<lib list ('' = product lib)>" "<debug list>"
LIBS=${1:-product}
DBG=${2:-0 2}
for lib in $LIBS; do
  for d in $DBG; do
    if [ "$lib" = product ]; then unset RVE_B200_LIB; else export RVE_B200_LIB=$PWD/$lib; fi
    RVE_CULL_DEBUG=$d python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$lib debug=$d kernel_ms', round(d['roofline']['kernel_ms'],4), 'frac', round(d['roofline']['frac'],3), 'visible', d['config']['visible'])"
  done
done
