#!/bin/bash
# N=2 step with parts of the cross-stream exchange protocol switched off (timing experiments; results of the exchange are not used)
run() { env $(echo $1 | tr , ' ') python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 40 --warmup 5 --no-e2e --no-c5 --no-gather --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$1', round(d['ms_per_step']*1e3,1), 'us step; cull', round(d['roofline']['kernel_ms']*1e3,1), 'by rank', [round(x*1e3,1) for x in d['ms_per_step_by_rank']])"; }
for v in $@; do run $v; done
