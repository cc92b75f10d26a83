"""Seam B2 from C++: tests/cpp/rollouts_shim.cc — the SamplingPlanner::Rollouts / OptimizePolicy bodies of
INTEGRATION.md §2, compiled with g++ against include/hmj_b200.h — must hand a C++ host the same bits the Python host gets."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIM = os.path.join(REPO, "tests", "cpp", "_build", "librollouts_shim.so")


def load_shim():
    if not os.path.exists(SHIM):
        subprocess.check_call(["make", "-C", os.path.join(REPO, "tests", "cpp")])
    lib = C.CDLL(SHIM)
    vp = C.c_void_p
    lib.shim_last_error.restype = C.c_char_p
    lib.shim_create.restype = vp
    lib.shim_create.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double]
    lib.shim_destroy.argtypes = [vp]
    lib.shim_set_state.argtypes = [vp, vp, C.c_double, vp, C.c_int]
    lib.shim_rollouts.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, vp, vp]
    lib.shim_optimize.argtypes = [vp, C.c_int, vp, vp, C.c_double, C.c_ulonglong, C.c_ulonglong, C.c_int, vp, vp]
    return lib


def test_shim_builds_and_links_the_cabi():
    """CPU-side: the C++ translation unit compiles against the public headers and resolves every hmj_* symbol it uses."""
    lib = load_shim()
    for sym in ("shim_create", "shim_rollouts", "shim_optimize"):
        assert hasattr(lib, sym)


@pytest.mark.gpu
def test_cpp_host_matches_python_host(oracle):
    import humanoid_mujoco_b200 as hmj
    from test_gpu_planner import settled_task_state

    lib = load_shim()
    state, time = settled_task_state(oracle)
    n, P, H = 32, 3, 18
    pl = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=2)
    pl.SetState(state[:19], state[19:], time)
    pl.UpdateNominalPolicy(H)
    rng = np.random.default_rng(4)
    scale = 0.5 * (pl.ctrlrange[:, 1] - pl.ctrlrange[:, 0])
    noise = rng.standard_normal((n, P, 12)) * scale * 0.05
    pl.Rollouts(H, noise=noise)
    times, values, want = pl.candidate_times.copy(), pl.candidate_values.copy(), pl.total_return.copy()
    nominal = pl.policy.plan.arrays()[1].copy()

    model = pl.batch.model
    h = lib.shim_create(model.ctypes.data, n, 0, P, pl.interpolation_, pl.timestep)
    assert h, lib.shim_last_error()
    st = np.ascontiguousarray(state)
    assert lib.shim_set_state(h, st.ctypes.data, time, times.ctypes.data, P) == 0
    ret, fail = np.zeros(n), np.zeros(n, dtype=np.int32)
    assert lib.shim_rollouts(h, n, H, values.ctypes.data, 1, ret.ctypes.data, fail.ctypes.data) == 0, lib.shim_last_error()
    assert np.array_equal(ret, want) and not fail.any()
    # OptimizePolicy through the C++ host: device candidates + built-in exchange (single process) == Python host
    pl.OptimizePolicyCandidates(1, H, noise=noise)
    out = np.zeros(3 + P * 12)
    nz = np.ascontiguousarray(noise)
    assert lib.shim_optimize(h, H, nominal.ctypes.data, nz.ctypes.data, 0.05, 2, 0, 0, None, out.ctypes.data) == 0, lib.shim_last_error()
    assert int(out[0]) == pl.winner and out[1] == pl.winner_return and out[2] == pl.nominal_return
    assert np.array_equal(out[3:].reshape(P, 12), pl._winner_values)
    # device-drawn noise: same (seed, iteration, global index) stream on both hosts
    pl2 = hmj.SamplingPlanner(num_trajectory=n, device=0, seed=9)
    pl2.SetState(state[:19], state[19:], time)
    pl2.policy.plan = pl.policy.plan.copy()
    pl2.policy.plan.times, pl2.policy.plan.values = list(times), [v.copy() for v in nominal]
    pl2.OptimizePolicyCandidates(1, H)
    assert lib.shim_optimize(h, H, nominal.ctypes.data, None, pl2.noise_exploration[0], 9, 0, 0, None, out.ctypes.data) == 0
    assert int(out[0]) == pl2.winner and out[1] == pl2.winner_return
    assert np.array_equal(out[3:].reshape(P, 12), pl2._winner_values)
    lib.shim_destroy(h)


@pytest.mark.gpu
def test_builtin_nccl_exchange_two_gpus():
    """hmj_comm_* + hmj_plan_exchange (the collective a C++ host uses) against the torch.distributed path; needs 2 GPUs."""
    import sys

    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs (run with gpurun --gpus 2)")
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                          "127.0.0.1", "--master-port", "29541", os.path.join(REPO, "tests", "multi_gpu", "exchange_nccl.py")],
                         capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "EXCHANGE OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
