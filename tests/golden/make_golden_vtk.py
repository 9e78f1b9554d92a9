#!/usr/bin/env python
"""Generates tests/golden/vtk/* with the UNMODIFIED reference writer (oracle/_ref/libref_writer.so = /root/reference/writer.cpp
behind C entry points, built by `make -C oracle ref`).  Run in the build container (needs /root/reference):
    python tests/golden/make_golden_vtk.py
State: the 150-particle aggregate after its sinter constructor (bond set + connectors from the oracle), with particle
orientations drawn from a fixed-seed generator so that the theta -> triad conversion of write_particles is exercised."""
import ctypes as C
import os
import shutil
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)


def vtk_state():
    """(x, theta, r_part, bonded [n, n] uint8, connectors [n, n, 15]) of the golden dump."""
    from libgran_b200 import scenarios as S
    from oracle import bindings as ob

    scn = S.aggregate()
    cfg = scn.cfg
    o = ob.OracleSystem(cfg, *scn.state)
    n = cfg.n
    rng = np.random.default_rng(11)
    theta = rng.uniform(-3.0, 3.0, size=(n, 3))
    theta[:5] = 0.0
    theta[5] = [1e-9, -2e-9, 3e-9]
    bonded = np.zeros((n, n), dtype=np.uint8)
    conn = np.zeros((n, n, 15))
    for i in range(n):
        for j in range(n):
            if i != j:
                b = o.bond(i, j)
                if b is not None:
                    bonded[i, j] = 1
                    conn[i, j] = b[0].reshape(15)
    return scn.x0, theta, cfg.r_part, bonded, conn


def reference_files(x, theta, r_part, bonded, conn, out_dir):
    lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref_writer.so"))
    dp = C.POINTER(C.c_double)
    n = len(x)
    x = np.ascontiguousarray(x, dtype=np.float64)
    theta = np.ascontiguousarray(theta, dtype=np.float64)
    bonded = np.ascontiguousarray(bonded, dtype=np.uint8)
    conn = np.ascontiguousarray(conn, dtype=np.float64)
    lib.refw_write_particles.argtypes = [C.c_char_p, C.c_int64, dp, dp, C.c_double]
    lib.refw_write_necks.argtypes = [C.c_char_p, C.c_int64, dp, C.c_double, C.c_void_p]
    lib.refw_write_spring_connectors.argtypes = [C.c_char_p, C.c_int64, C.c_double, dp, C.c_void_p]
    d = os.fsencode(out_dir)
    lib.refw_write_particles(d, n, x.ctypes.data_as(dp), theta.ctypes.data_as(dp), r_part)
    lib.refw_write_necks(d, n, x.ctypes.data_as(dp), r_part, bonded.ctypes.data)
    lib.refw_write_spring_connectors(d, n, r_part, conn.ctypes.data_as(dp), bonded.ctypes.data)


if __name__ == "__main__":
    out = os.path.join(HERE, "vtk")
    os.makedirs(out, exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp:
        reference_files(*vtk_state(), tmp)
        for name in sorted(os.listdir(tmp)):  # numbered _1 by the reference's static counters
            shutil.copy(os.path.join(tmp, name), os.path.join(out, name.replace("_1.vtk", ".vtk")))
            print(name, os.path.getsize(os.path.join(tmp, name)))
