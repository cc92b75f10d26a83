#!/bin/bash
# usage: ab_env.sh VAR val1 val2 ... : bench (no planner, no CPU baseline) once per value of the environment variable
VAR=$1; shift
for v in "$@"; do
  export $VAR="$v"
  python bench.py --steps 200 --warmup 10 --no-plan --no-cpu 2>/dev/null > gpurun_out/ab_env.json
  python -c "
import json; d=json.load(open('gpurun_out/ab_env.json')); print('$VAR=$v', '%.4g' % d['value'], '%.4g' % d['e2e']['value'], {k: round(x, 4) for k, x in d['roofline']['kernel_ms'].items()})"
done
