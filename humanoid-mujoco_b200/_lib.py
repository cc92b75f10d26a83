"""ctypes loader of libhmj_b200.so.  Fails loudly: no library, no engine."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class HmjError(RuntimeError):
    pass


def lib_path() -> str:
    # HMJ_B200_LIB: developer override to A/B-test an alternative build of the same engine
    return os.environ.get("HMJ_B200_LIB") or os.path.join(_HERE, "lib", "libhmj_b200.so")


def load_library() -> C.CDLL:
    """Load the CUDA engine.  Raises HmjError if the shared library has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise HmjError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(path)
    vp, ip, dp = C.c_void_p, C.POINTER(C.c_int32), C.c_void_p
    lib.hmj_last_error.restype = C.c_char_p
    lib.hmj_batch_create.argtypes = [vp, C.c_int, C.c_int, C.POINTER(vp)]
    lib.hmj_batch_destroy.argtypes = [vp]
    lib.hmj_batch_destroy.restype = None
    lib.hmj_batch_size.argtypes = [vp]
    lib.hmj_batch_sync.argtypes = [vp]
    lib.hmj_default_params.argtypes = [vp, vp]
    lib.hmj_batch_set_params.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.hmj_batch_get_params.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.hmj_batch_set_state.argtypes = [vp, dp, dp, dp, dp]
    lib.hmj_batch_get_state.argtypes = [vp, dp, dp, dp, dp]
    lib.hmj_batch_set_ctrl.argtypes = [vp, dp]
    lib.hmj_batch_set_xfrc.argtypes = [vp, vp, dp]
    lib.hmj_batch_step.argtypes = [vp, dp, C.c_int]
    lib.hmj_batch_step_dev.argtypes = [vp, dp, C.c_int]
    lib.hmj_batch_forward.argtypes = [vp]
    lib.hmj_batch_run_stage.argtypes = [vp, C.c_int]
    lib.hmj_batch_get_contacts.argtypes = [vp, vp, vp]
    lib.hmj_batch_get_array.argtypes = [vp, C.c_char_p, vp, C.c_size_t]
    lib.hmj_batch_set_array.argtypes = [vp, C.c_char_p, vp, C.c_size_t]
    lib.hmj_batch_device_ptr.argtypes = [vp, C.c_char_p]
    lib.hmj_batch_device_ptr.restype = vp
    lib.hmj_batch_stream.argtypes = [vp]
    lib.hmj_batch_stream.restype = vp
    lib.hmj_batch_kernel_launches.argtypes = [vp]
    lib.hmj_support_cells.argtypes = [vp, vp, C.c_int, vp, C.c_int, vp, vp]
    lib.hmj_env_obs_reward.argtypes = [vp, dp, dp, dp, C.c_double, dp, dp, vp, vp, vp]
    lib.hmj_env_obs_reward_dev.argtypes = [vp, dp, dp, dp, C.c_double, dp, dp, vp, vp, vp]
    lib.hmj_env_obs_reward_noalloc.argtypes = [vp, dp, dp, dp, C.c_double, dp, dp, vp, vp, vp]
    lib.hmj_env_obs_reward_noalloc_max_envs.argtypes = []
    lib.hmj_env_step.argtypes = [vp, dp, vp, dp, dp, dp, dp, C.c_double, C.c_int, dp, dp, vp, vp, vp, dp, dp]
    lib.hmj_env_io.argtypes = [vp] + [C.POINTER(vp)] * 9
    lib.hmj_env_step_fused.argtypes = [vp, C.c_int, C.c_int, C.c_double, C.c_double, C.c_ulonglong, C.c_int]
    lib.hmj_env_step_fused_from.argtypes = [vp, vp, C.c_int, C.c_double, C.c_double, C.c_ulonglong, C.c_int]
    lib.hmj_pinned_alloc.argtypes = [C.c_size_t, C.POINTER(vp)]
    lib.hmj_pinned_free.argtypes = [vp]
    lib.hmj_batch_profile.argtypes = [vp, C.c_int]
    lib.hmj_batch_profile_read.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double),
                                           C.POINTER(C.c_int)]
    lib.hmj_fp64_peak.argtypes = [C.c_int, C.POINTER(C.c_double)]
    lib.hmj_plan_set_task.argtypes = [vp, C.c_int, vp, vp, dp, dp, C.c_double, C.c_double]
    lib.hmj_plan_rollouts.argtypes = [vp, dp, C.c_double, dp, dp, C.c_int, C.c_int, C.c_int, C.c_int, dp, vp]
    lib.hmj_plan_rollouts_noisy.argtypes = [vp, dp, C.c_double, dp, dp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, dp,
                                            C.c_ulonglong, C.c_int, dp, vp]
    lib.hmj_plan_rollouts_dev.argtypes = [vp, dp, C.c_double, dp, dp, C.c_int, C.c_int, C.c_int, dp]
    lib.hmj_plan_argmin.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_int32)]
    lib.hmj_plan_fetch.argtypes = [vp, C.c_int, C.c_int, dp, dp, dp, dp, dp, dp]
    lib.hmj_env_pool_config.argtypes = [vp, C.c_int, C.c_double, C.c_double, C.c_ulonglong, C.c_double]
    lib.hmj_env_reset.argtypes = [vp, vp, C.c_int]
    lib.hmj_env_reset_io.argtypes = [vp] + [C.POINTER(vp)] * 3
    lib.hmj_batch_spare_rows.argtypes = [vp]
    lib.hmj_env_terminal.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(C.c_int32)]
    lib.hmj_env_pool_stats.argtypes = [vp, vp]
    lib.hmj_plan_iterate.argtypes = [vp, dp, C.c_double, dp, dp, C.c_int, C.c_int, C.c_int, dp, C.c_double, C.c_double,
                                     C.c_ulonglong, C.c_ulonglong, C.c_int, vp]
    lib.hmj_plan_iterate_ex.argtypes = [vp, dp, C.c_double, dp, dp, C.c_int, C.c_int, C.c_int, dp, C.c_double, C.c_double, dp,
                                        C.c_ulonglong, C.c_ulonglong, C.c_int, vp]
    lib.hmj_batch_inverse.argtypes = [vp, dp, dp]
    lib.hmj_fd_transition.argtypes = [vp, C.c_int, dp, dp, dp, dp, C.c_double, C.c_int, dp, dp, dp, dp]
    lib.hmj_plan_record_size.argtypes = [C.c_int]
    lib.hmj_plan_record_ptr.argtypes = [vp]
    lib.hmj_plan_record_ptr.restype = vp
    lib.hmj_plan_select.argtypes = [vp, vp, C.c_int, C.c_int, dp]
    lib.hmj_plan_population.argtypes = [vp, C.c_int, dp, vp, dp]
    lib.hmj_batch_host_launches.argtypes = [vp]
    _LIB = lib
    return lib


def check(rc: int):
    if rc != 0:
        raise HmjError(f"hmj_b200 error {rc}: {load_library().hmj_last_error().decode()}")
