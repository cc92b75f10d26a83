#!/bin/bash
# A/B of two builds of the library on the bench workload: lib/ (A) vs lib_prof/ (B)
for v in A B; do
  if [ $v = B ]; then export HMJ_B200_LIB=humanoid-mujoco_b200/lib_prof/libhmj_b200.so; fi
  python bench.py --steps 200 --warmup 10 --no-plan 2>/dev/null > gpurun_out/ab_$v.json
  python -c "
import json; d=json.load(open('gpurun_out/ab_$v.json')); print('$v', '%.4g' % d['value'], '%.4g' % d['e2e']['value'], {k: round(x, 4) for k, x in d['roofline']['kernel_ms'].items()})"
done
