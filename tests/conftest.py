import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for sub in ("", "tools", "oracle"):
    p = os.path.join(REPO, sub)
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle_py import Oracle

    return Oracle()


def standup_state(model, rng=None, rf=0.0):
    """CPUEnv.reset start pose (simulation/cpu_env.py:282-328, standup reward selected)."""
    import numpy as np

    q = np.array(model["qpos0"][0], dtype=np.float64)
    q[2] = -0.6
    q[3:7] = [-0.5, -0.5, 0.5, 0.5]
    if rng is not None and rf > 0:
        q[7:] += rng.uniform(-0.2 * rf, 0.2 * rf, 12)
        q[2] += rng.uniform(0, 0.1 * rf)
        q[3:7] += rng.uniform(-0.1 * rf, 0.1 * rf, 4)
    return q
