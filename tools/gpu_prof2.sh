#!/bin/bash
# ncu evidence at the STEADY STATE of both bench regimes (tools/dev/prof_steps.py): launch list + full-set capture of the
# step kernels of the last steps.  One pipe, direct launches, so the list reads as one step after the other.
# usage: tools/gpu_prof2.sh <tag>
TAG=${1:-prof}
mkdir -p gpurun_out
export HMJ_PIPES=1 HMJ_GRAPH=0
for W in "C3 1.0 150" "settled 0.02 50"; do
  set -- $W
  NAME=$1; SCALE=$2; LEAD=$3
  CMD="python tools/dev/prof_steps.py $SCALE $LEAD 3"
  $CMD > gpurun_out/${TAG}_${NAME}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_${NAME}_plain.log; exit 1; }
  tail -1 gpurun_out/${TAG}_${NAME}_plain.log
  SKIP=$(python - <<PY
import re
t=open("gpurun_out/${TAG}_${NAME}_plain.log").read()
print(int(re.search(r"steps: (\d+)", t).group(1)) * 10)
PY
)
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'^k_' -s $SKIP -c 30 --csv --log-file gpurun_out/${TAG}_${NAME}_launches.csv $CMD > gpurun_out/${TAG}_${NAME}_ncu1.log 2>&1
  ncu --set full --clock-control none --import-source on -k regex:'^k_(kin|dyn|factor|broad|narrow|classify|solve)' -s $SKIP -c 10 -o gpurun_out/${TAG}_${NAME}_step $CMD > gpurun_out/${TAG}_${NAME}_ncu2.log 2>&1
done
ls -la gpurun_out | grep ${TAG}
