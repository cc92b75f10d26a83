#!/usr/bin/env python
"""Generate golden vectors for the mjpc spline + norm from the REFERENCE'S OWN sources.

Runs in the build container only (needs /root/reference): `make -C oracle ref` compiles
mujoco_mpc/mjpc/norm.cc and mjpc/spline/spline.cc in place into oracle/_ref/libmjpc_ref.so;
this script drives that library on seeded inputs and commits inputs + outputs to
tests/golden/mjpc_golden.npz so the pins travel to machines without the reference tree.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))


def main():
    subprocess.check_call(["make", "-C", os.path.join(REPO, "oracle"), "ref"])
    lib = C.CDLL(os.path.join(REPO, "oracle", "_ref", "libmjpc_ref.so"))
    lib.ref_norm.restype = C.c_double
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    rng = np.random.default_rng(20241104)
    out = {}
    # ---- splines: random node counts / dims / query times (inside, outside, on nodes), 3 interpolations
    cases = []
    for n in (1, 2, 3, 4, 7, 12):
        for dim in (1, 2, 12):
            times = np.cumsum(rng.uniform(0.01, 0.5, n)) + rng.uniform(-1, 1)
            values = rng.uniform(-1, 1, (n, dim))
            q = np.concatenate([rng.uniform(times[0] - 0.3, times[-1] + 0.3, 12), times])
            cases.append((times, values, q))
    sp_t, sp_v, sp_q, sp_o, sp_meta = [], [], [], [], []
    for times, values, q in cases:
        for interp in (0, 1, 2):
            for t in q:
                o = np.zeros(values.shape[1])
                lib.ref_spline_sample(p(times), p(values), len(times), values.shape[1], interp, C.c_double(t), p(o))
                sp_meta.append((len(times), values.shape[1], interp))
                sp_t.append(np.pad(times, (0, 12 - len(times))))
                sp_v.append(np.pad(values, ((0, 12 - values.shape[0]), (0, 12 - values.shape[1]))))
                sp_q.append(t)
                sp_o.append(np.pad(o, (0, 12 - o.size)))
    out.update(spline_meta=np.array(sp_meta), spline_times=np.array(sp_t), spline_values=np.array(sp_v),
               spline_query=np.array(sp_q), spline_out=np.array(sp_o))
    # ---- norms: all 8 types of norm.h:24-35 with the parameter sets of norm_test.cc
    ntypes = [(0, []), (1, [0.1, 2.0]), (2, [0.1]), (3, [0.3]), (5, [1.5]), (6, [0.1]), (7, [0.1, 2.0]), (8, [0.2])]
    nm, nx, no = [], [], []
    for ntype, prm in ntypes:
        for n in (1, 2, 12):
            for _ in range(8):
                x = rng.uniform(-2, 2, n)
                prm3 = np.zeros(3)
                prm3[: len(prm)] = prm
                y = lib.ref_norm(p(x), p(prm3), n, ntype)
                nm.append((ntype, n))
                nx.append(np.concatenate([prm3, np.pad(x, (0, 12 - n))]))
                no.append(y)
    out.update(norm_meta=np.array(nm), norm_in=np.array(nx), norm_out=np.array(no))
    np.savez_compressed(os.path.join(HERE, "mjpc_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "mjpc_golden.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
