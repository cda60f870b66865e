#!/bin/bash
# One gpurun call on an N-GPU box: the round's driver-shaped records (GPU tests at N=1, reference arm, bench at the given N).
#   gpurun --gpus N -- bash tools/record_round.sh N TAG
N=${1:-1}; TAG=${2:-r02c}
if [ "$N" = "1" ]; then
  python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_gpu_tests.log 2>&1; tail -2 gpurun_out/${TAG}_gpu_tests.log
  python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${TAG}_smoke.log 2>&1; tail -1 gpurun_out/${TAG}_smoke.log
  python bench.py --impl reference --steps 20 --warmup 5 2>/dev/null | grep "^{" > gpurun_out/${TAG}_bench_reference_arm.json
  python bench.py --steps 20 --warmup 5 2> gpurun_out/${TAG}_bench_n1.err | grep "^{" > gpurun_out/${TAG}_bench_n1.json; tail -2 gpurun_out/${TAG}_bench_n1.err
else
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 295$N$N bench.py --gpus $N --steps 20 --warmup 5 \
      2> gpurun_out/${TAG}_bench_n$N.err | grep "^{" > gpurun_out/${TAG}_bench_n$N.json; tail -2 gpurun_out/${TAG}_bench_n$N.err | cut -c1-200
fi
ls -la gpurun_out/${TAG}_* | awk '{print $5, $9}'
