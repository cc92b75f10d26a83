#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (both arms), then ncu launch lists + full captures at the steady state of
# both bench regimes (tools/gpu_prof2.sh).
# usage: tools/gpu_round.sh <tag> [noprof]
TAG=${1:-round}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/${TAG}_gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${TAG}_pytest_gpu.log
tail -5 gpurun_out/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/${TAG}_smoke.log 2>&1; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py --steps 200 --warmup 10 > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench rc=$?"
timeout 900 python bench.py --steps 20 --warmup 5 --no-plan --no-c5 --no-cpu > gpurun_out/${TAG}_bench20.json 2>> gpurun_out/${TAG}_bench.err; echo "bench20 rc=$?"
timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref.json 2>> gpurun_out/${TAG}_bench.err
cut -c1-400 gpurun_out/${TAG}_bench.json; cut -c1-300 gpurun_out/${TAG}_bench_ref.json
if [ "$2" != "noprof" ]; then bash tools/gpu_prof2.sh ${TAG}; fi
