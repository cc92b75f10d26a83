#!/bin/bash
# Scaling record on one box with 8 GPUs: the driver's command line (torchrun, 20 timed steps, 5 warm-up) at 2, 4 and 8 ranks,
# plus the 2-GPU tests.  usage (inside gpurun --gpus 8): tools/gpu_scale.sh <tag>
TAG=${1:-scale}
mkdir -p gpurun_out
timeout 600 python -m pytest -m gpu -x -q tests/test_gpu_cpp_shim.py > gpurun_out/${TAG}_pytest_multigpu.log 2>&1; tail -2 gpurun_out/${TAG}_pytest_multigpu.log
NGPU=$(nvidia-smi -L | wc -l)
for N in 2 4 8; do
  if [ $N -gt $NGPU ]; then continue; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_n${N}.json 2> gpurun_out/${TAG}_bench_n${N}.err
  echo "N=$N rc=$?"; cut -c1-300 gpurun_out/${TAG}_bench_n${N}.json
done
