"""Single source of truth for binary layouts: parse the C headers into numpy dtypes.

``include/hmj_model.h`` (hmj_model_t, hmj_params_t, hmj_contact_t) and, for the tests,
``oracle/hmj_oracle.h`` (orc_data_t) are parsed into numpy structured dtypes with C
alignment rules (``align=True``) so that Python (compiler, tests, host wrappers) and
C/CUDA (oracle, kernels) agree byte for byte.
"""
from __future__ import annotations

import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(_HERE)
HEADER = os.path.join(REPO, "include", "hmj_model.h")

_CTYPES = {"int32_t": np.int32, "double": np.float64, "float": np.float32}


def parse_header(path: str, defines: dict | None = None, types: dict | None = None):
    """Return (structs: name -> dtype, defines: name -> int) for every typedef struct in `path`."""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    defines = dict(defines or {})
    types = dict(types or _CTYPES)
    for m in re.finditer(r"#define\s+((?:HMJ|ORC)_\w+)\s+(.+)", src):
        name, expr = m.group(1), m.group(2).strip()
        try:
            defines[name] = int(eval(expr, {}, dict(defines)))  # noqa: S307 (own header)
        except Exception:
            pass
    structs = {}
    for sm in re.finditer(r"typedef struct \w+ \{(.*?)\} (\w+);", src, re.S):
        body, sname = sm.group(1), sm.group(2)
        fields = []
        for stmt in body.split(";"):
            stmt = stmt.strip()
            if not stmt:
                continue
            m = re.match(r"(\w+)\s+(.*)", stmt, re.S)
            ctype, decls = m.group(1), m.group(2)
            if ctype not in types:
                raise ValueError(f"unknown type {ctype!r} in {sname}: {stmt!r}")
            for d in decls.split(","):
                d = d.strip()
                dm = re.match(r"(\w+)((?:\[[^\]]+\])*)$", d)
                name = dm.group(1)
                dims = tuple(
                    int(eval(x, {}, dict(defines)))  # noqa: S307
                    for x in re.findall(r"\[([^\]]+)\]", dm.group(2))
                )
                fields.append((name, types[ctype], dims) if dims else (name, types[ctype]))
        dt = np.dtype(fields, align=True)
        structs[sname] = dt
        types[sname] = dt
    return structs, defines, types


_STRUCTS, DEFINES, _TYPES = parse_header(HEADER)
MODEL_DTYPE = _STRUCTS["hmj_model_t"]
PARAMS_DTYPE = _STRUCTS["hmj_params_t"]
CONTACT_DTYPE = _STRUCTS["hmj_contact_t"]

DEFAULT_MODEL_PATH = os.path.join(REPO, "humanoid-mujoco_b200", "assets", "humanoid_model.bin")


def oracle_dtypes():
    """dtypes of the oracle's structs (tests only)."""
    structs, defines, _ = parse_header(os.path.join(REPO, "oracle", "hmj_oracle.h"), DEFINES, _TYPES)
    return structs, defines


def load_model(path: str = DEFAULT_MODEL_PATH) -> np.ndarray:
    """Load a blob written by compile_model.py; returns a 1-element structured array."""
    raw = np.fromfile(path, dtype=np.uint8)
    if raw.size != MODEL_DTYPE.itemsize:
        raise ValueError(f"model blob size {raw.size} != sizeof(hmj_model_t) {MODEL_DTYPE.itemsize}")
    m = raw.view(MODEL_DTYPE)
    if int(m["magic"][0]) != DEFINES["HMJ_MAGIC"] or int(m["version"][0]) != DEFINES["HMJ_VERSION"]:
        raise ValueError("model blob magic/version mismatch")
    return m
