#!/bin/bash
# development aid: time the C2 cull kernel with parts of it disabled (RVE_CULL_DEBUG bits: 1 no Hi-Z, 2 no phase B, 4 no phase A)
for rep in 1 2; do
for d in ${DEBUG_SET:-0 1 2 6}; do
  RVE_CULL_DEBUG=$d python bench.py --steps 20 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('debug=$d kernel_ms', round(d['roofline']['kernel_ms'],4), 'step_ms', round(d['ms_per_step'],4), 'visible', d['config']['visible'], d['clocks'])"
done
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.mem,power.draw,temperature.gpu --format=csv
