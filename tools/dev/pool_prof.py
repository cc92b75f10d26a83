"""Developer tool: what the episode pool costs per control step (16384 envs, rf = 0)."""
import ctypes as C, os, sys, time
import numpy as np
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))]
import humanoid_mujoco_b200 as hmj

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
acts = hmj.pinned_empty((16, n, 12)); acts[...] = np.random.default_rng(0).uniform(-1, 1, acts.shape)
for name, auto, limit, stag in (("auto off", False, None, 0), ("auto on, none ending", True, None, 0), ("auto on, 1%/step", True, 0.2, 100),
                                ("auto on, 0.25%/step", True, 0.8, 400)):
    e = hmj.BatchedHumanoidEnv(n, randomization_factor=0.0, max_simulation_time_override=limit, seed=1)
    e.auto_reset = auto
    e.reset()
    if stag:
        t = e.batch.get("time"); t += (np.arange(n) % stag) * 0.002; e.batch.set("time", t)
    for s in range(stag + 20):
        e.step(acts[s % 16])
    st0 = (C.c_longlong * 4)(); e.batch.lib.hmj_env_pool_stats(e.batch._h, st0)
    k, cnt = 200, 0
    t0 = time.perf_counter()
    for s in range(k):
        _o, _r, term, trunc, _i = e.step(acts[s % 16])
        if term is not e._all_false or trunc is not e._all_false:
            cnt += int((term | trunc).sum())
    e.batch.sync()
    dt = time.perf_counter() - t0
    st1 = (C.c_longlong * 4)(); e.batch.lib.hmj_env_pool_stats(e.batch._h, st1)
    print(f"{name:24s}: {1e3*dt/k:7.3f} ms/step  endings/step {cnt/k:7.2f}  rounds {st1[0]-st0[0]} drains {st1[1]-st0[1]} drain iters {st1[2]-st0[2]}"
          f"  ready left {e.batch.env_reset_io()['counts'][3]}", flush=True)
    e.batch.profile(True)
    for s in range(50):
        e.step(acts[s % 16])
    pr = e.batch.profile_read(); e.batch.profile(False)
    print(f"    physics kernels (events, direct launches): smooth {1e3*pr['smooth_ms']/pr['steps']:.0f} collide {1e3*pr['collide_ms']/pr['steps']:.0f} solve {1e3*pr['solve_ms']/pr['steps']:.0f} us;"
          f" mean ncon {e.batch.get('ncon').mean():.2f} nefc {e.batch.get('nefc').mean():.1f} niter {e.batch.get('niter').mean():.2f}", flush=True)
    e.close()
