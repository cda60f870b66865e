#!/bin/bash
# development aid: C2 step under env-var variants, one line each
run() { echo -n "$1: "; env $1 python bench.py --steps 20 --warmup 5 --no-e2e --no-c5 --no-stages --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print(round(d['ms_per_step']*1e3,1), 'us step; cull', round(d['roofline']['kernel_ms']*1e3,1), 'pyr', round(d['roofline']['pyramid_kernel_ms']*1e3,1))
"; }
for v in "$@"; do run "$v"; done
