#!/bin/bash
# pyramid build A/B on one box: persistent launch, its phases switched off one by one (development), the two-launch build
B="python bench.py --steps 20 --warmup 5 --no-e2e --no-c5 --no-stages --no-cpu-baseline"
run() { env $(echo $1 | tr , ' ') $B 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step']*1e3,1), 'us step; cull', round(d['roofline']['kernel_ms']*1e3,1), 'pyr', round(d['roofline']['pyramid_kernel_ms']*1e3,2))"; }
for v in $@; do run $v; done
