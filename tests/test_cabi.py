"""CPU-side checks of the boundary: the shared library loads and exports every symbol that
include/hmj_b200.h declares (no compute calls: there is no GPU here), and refuses to run without one."""
import ctypes as C
import os
import re

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(REPO, "include", "hmj_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hmj_\w+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import humanoid_mujoco_b200 as hmj

    if not os.path.exists(hmj.lib_path()):
        import __graft_entry__ as g

        g.build()
    lib = C.CDLL(hmj.lib_path())
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/hmj_b200.h but not exported"


def test_no_cpu_fallback():
    import numpy as np

    import humanoid_mujoco_b200 as hmj
    from conftest import gpu_available

    if gpu_available():
        pytest.skip("GPU present")
    with pytest.raises(hmj.HmjError):
        hmj.HumanoidBatch(4)
    # the product package must not import the oracle
    import sys

    pkg_dir = os.path.join(REPO, "humanoid-mujoco_b200")
    for root, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                txt = open(os.path.join(root, f)).read()
                assert "oracle_py" not in txt and "libhmj_oracle" not in txt, f
                assert not re.search(r'#include\s*[<"][^>"]*oracle', txt), f
    assert np is not None and sys is not None


def test_model_blob_layout_matches_header():
    import humanoid_mujoco_b200 as hmj
    from model_layout import MODEL_DTYPE, PARAMS_DTYPE
    from humanoid_mujoco_b200.batch import MODEL_DTYPE as M2, PARAMS_DTYPE as P2

    assert MODEL_DTYPE == M2 and PARAMS_DTYPE == P2
    m = hmj.load_model()
    assert int(m["npair"][0]) == 37 and int(m["task_nresidual"][0]) == 28
