#!/bin/bash
# Developer A/B: quick bench (both regimes, value + e2e) under a list of environment settings.
# usage: tools/dev/ab.sh <tag> "VAR=1 VAR2=2" "..." ...     (an empty string = defaults)
TAG=$1; shift
mkdir -p gpurun_out
i=0
for cfg in "$@"; do
  out=gpurun_out/${TAG}_ab${i}.json
  env $cfg timeout 600 python bench.py --steps 200 --warmup 10 --no-plan --no-c5 --no-cpu > $out 2> gpurun_out/${TAG}_ab${i}.err
  python - "$cfg" $out <<'P'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    w = d["workloads"]
    print("[%s]" % sys.argv[1], " ".join("%s value %.4g e2e %.4g" % (k, v["value"], v["e2e"]["value"]) for k, v in w.items()))
except Exception as e:
    print("[%s] failed: %s" % (sys.argv[1], e))
P
  i=$((i+1))
done
