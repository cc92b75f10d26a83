#!/bin/bash
# ncu evidence for one round: launch list of a short bench run + full-set capture of the step kernels.
# usage: tools/gpu_prof.sh <tag>   (run on the GPU box through gpurun; outputs under gpurun_out/)
# One pipe and direct launches (HMJ_PIPES=1 HMJ_GRAPH=0) so that the list reads as one step after the other:
# 8 kernels per step (smooth, factor, broad, narrow, classify, solve x 3) + the obs/reward epilogue.
TAG=${1:-prof}
mkdir -p gpurun_out
export HMJ_PIPES=1 HMJ_GRAPH=0
CMD="python bench.py --steps 3 --warmup 3 --no-cpu --no-plan"
$CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -s 2427 -c 108 --csv --log-file gpurun_out/${TAG}_launches.csv $CMD > gpurun_out/${TAG}_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'^k_(smooth|factor|broad|narrow|classify|solve)' -s 2454 -c 16 -o gpurun_out/${TAG}_step $CMD > gpurun_out/${TAG}_ncu2.log 2>&1
ls -la gpurun_out
