for v in "" c3 c4; do
  if [ -n "$v" ]; then export HMJ_B200_LIB=$PWD/humanoid-mujoco_b200/lib/alt/$v.so; fi
  python bench.py --steps 200 --warmup 10 --no-cpu 2> gpurun_out/bench.err | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$v', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'])"
done
