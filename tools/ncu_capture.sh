#!/bin/bash
# One gpurun call: the bench's launch list (gpu__time_duration per launch) and one `ncu --set full` capture of every hot kernel.
# Each command runs plain first (B200_PROFILING.md); outputs under gpurun_out/.   usage: tools/ncu_capture.sh [TAG]
set -x
TAG=${1:-r02}
B="python bench.py --steps 2 --warmup 3 --no-e2e --no-c5 --no-stages --no-cpu-baseline"
$B > gpurun_out/ncu_plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${TAG}_launches.csv $B > gpurun_out/ncu_bench.log 2>&1
python tools/ncu_targets.py > gpurun_out/ncu_plain_targets.log 2>&1 && \
ncu --set full --clock-control none --import-source on \
    -k regex:'cull_kernel|cull_views_kernel|pyramid_persistent|pyramid_build_tma_kernel|pyramid_upper_kernel|skin_kernel|cluster_assign_kernel|full_rows_kernel' \
    -c 40 -f -o gpurun_out/${TAG}_prof python tools/ncu_targets.py > gpurun_out/ncu_targets.log 2>&1
ls -la gpurun_out/ | tail -8
tail -3 gpurun_out/ncu_targets.log
