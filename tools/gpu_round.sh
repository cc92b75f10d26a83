#!/bin/bash
# One GPU-box visit: parity tests, bench (both arms), ncu launch list + full capture of the step kernels.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python bench.py --steps 200 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json
python bench.py --impl reference --steps 50 --warmup 5 > gpurun_out/bench_ref.json 2>> gpurun_out/bench.err
CMD="python bench.py --steps 3 --warmup 3 --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 60 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_(smooth|collide|solve)' -s 915 -c 6 -o gpurun_out/prof_step $CMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
