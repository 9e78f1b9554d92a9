"""ctypes bindings of the two CPU checkers under ``oracle/``.

TEST INFRASTRUCTURE ONLY.  Only ``tests/``, ``__graft_entry__.smoke()`` and the CPU-baseline /
reference arm of ``bench.py`` may import this module; the product package ``libgran_b200`` never
does (it fails loudly when its CUDA library is missing instead of falling back to anything here).

* ``RefSystem``    -- ``oracle/_ref/libref_oracle.so``: the UNMODIFIED reference headers driven in
  lock-step (``oracle/ref_harness.cpp``), parity build (``-ffp-contract=off``).
  ``RefSystem(..., timing=True)`` loads ``libref_timing.so`` (the reference's own -O3 flags).
* ``OracleSystem`` -- ``oracle/libdem_oracle.so``: the sparse-state CPU restatement
  (``oracle/dem_oracle.cpp``), usable at sizes where the reference's 72..232 B * N^2 tables do not fit.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

MODEL_CONTACT, MODEL_HAMAKER, MODEL_SINTER, MODEL_FACET, MODEL_ALT_SINTER = 1, 2, 4, 8, 16
FIELD_X, FIELD_V, FIELD_A, FIELD_THETA, FIELD_OMEGA, FIELD_ALPHA = range(6)
FIELDS = {"x": 0, "v": 1, "a": 2, "theta": 3, "omega": 4, "alpha": 5}


class DemConfigStruct(C.Structure):
    """Mirror of ``struct dem_config`` (oracle/dem_config.h)."""

    _fields_ = [
        ("n", C.c_int64),
        ("r_part", C.c_double),
        ("mass", C.c_double),
        ("inertia", C.c_double),
        ("model_mask", C.c_int32),
        ("use_neighbor_list", C.c_int32),
        ("r_verlet", C.c_double),
        ("contact", C.c_double * 15),
        ("hamaker_A", C.c_double),
        ("hamaker_h0", C.c_double),
        ("k_d", C.c_double),
        ("gamma_d", C.c_double),
        ("critical_separation", C.c_double),
        ("rng_seed", C.c_uint64),
        ("facet_vertices", C.c_double * 9),
        ("facet_v", C.c_double * 3),
        ("surf_contact", C.c_double * 15),
        ("surf_A", C.c_double),
        ("surf_h0", C.c_double),
        ("alt_bond", C.c_double * 8),
        ("hamaker_kind", C.c_int32),
        ("integrator", C.c_int32),
        ("hamaker_cutoff", C.c_double),
        ("fixed_particle", C.c_int64),
    ]


def to_struct(cfg) -> DemConfigStruct:
    """Convert a ``libgran_b200.scenarios.DemConfig``-like object (duck-typed) to the C struct."""
    s = DemConfigStruct()
    s.n = int(cfg.n)
    s.r_part, s.mass, s.inertia = float(cfg.r_part), float(cfg.mass), float(cfg.inertia)
    s.model_mask = int(cfg.model_mask)
    s.use_neighbor_list = int(bool(cfg.use_neighbor_list))
    s.r_verlet = float(cfg.r_verlet)
    for k, val in enumerate(cfg.contact):
        s.contact[k] = float(val)
    s.hamaker_A, s.hamaker_h0 = float(cfg.hamaker_A), float(cfg.hamaker_h0)
    s.k_d, s.gamma_d, s.critical_separation = float(cfg.k_d), float(cfg.gamma_d), float(cfg.critical_separation)
    s.rng_seed = int(cfg.rng_seed)
    for k, val in enumerate(np.asarray(cfg.facet_vertices, dtype=np.float64).reshape(-1)):
        s.facet_vertices[k] = float(val)
    for k, val in enumerate(cfg.facet_v):
        s.facet_v[k] = float(val)
    for k, val in enumerate(cfg.surf_contact):
        s.surf_contact[k] = float(val)
    s.surf_A, s.surf_h0 = float(cfg.surf_A), float(cfg.surf_h0)
    for k, val in enumerate(getattr(cfg, "alt_bond", (0.0,) * 8)):
        s.alt_bond[k] = float(val)
    s.hamaker_kind = int(getattr(cfg, "hamaker_kind", 0))
    s.integrator = int(getattr(cfg, "integrator", 0))
    s.hamaker_cutoff = float(getattr(cfg, "hamaker_cutoff", 0.0))
    s.fixed_particle = int(getattr(cfg, "fixed_particle", -1))
    return s


def _dp(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def _c3(a, n) -> np.ndarray:
    out = np.ascontiguousarray(np.asarray(a, dtype=np.float64).reshape(n, 3))
    return out


# ----------------------------------------------------------------------------------------------
# oracle/_ref : the reference itself
# ----------------------------------------------------------------------------------------------
_REF_LIBS = {}


def ref_lib_path(timing: bool = False) -> str:
    return os.path.join(_HERE, "_ref", "libref_timing.so" if timing else "libref_oracle.so")


def ref_available(timing: bool = False) -> bool:
    return os.path.exists(ref_lib_path(timing))


def load_ref(timing: bool = False):
    key = bool(timing)
    if key in _REF_LIBS:
        return _REF_LIBS[key]
    lib = C.CDLL(ref_lib_path(timing))
    dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_void_p
    lib.ref_create.restype = vp
    lib.ref_create.argtypes = [C.POINTER(DemConfigStruct), dp, dp, dp, dp]
    lib.ref_destroy.argtypes = [vp]
    lib.ref_update_neighbor_list.argtypes = [vp]
    lib.ref_do_step.argtypes = [vp, C.c_double]
    lib.ref_run.argtypes = [vp, C.c_double, C.c_int64, C.c_int64]
    lib.ref_facet_update_positions.argtypes = [vp, C.c_double]
    lib.ref_get_field.argtypes = [vp, C.c_int, dp]
    lib.ref_n_pairs.restype = C.c_int64
    lib.ref_n_pairs.argtypes = [vp]
    lib.ref_get_rows.argtypes = [vp, ip, ip]
    lib.ref_get_contact_springs.argtypes = [vp, C.c_int64, C.c_int64, dp]
    lib.ref_get_bond.restype = C.c_int
    lib.ref_get_bond.argtypes = [vp, C.c_int64, C.c_int64, dp, dp]
    lib.ref_get_facet_vertices.argtypes = [vp, dp]
    lib.ref_total_kinetic_energy.restype = C.c_double
    lib.ref_total_kinetic_energy.argtypes = [dp, dp, C.c_int64, C.c_double, C.c_double]
    lib.ref_linear_momentum.restype = C.c_double
    lib.ref_linear_momentum.argtypes = [dp, C.c_int64, C.c_double]
    lib.ref_angular_momentum.restype = C.c_double
    lib.ref_angular_momentum.argtypes = [dp, dp, dp, C.c_int64, C.c_double, C.c_double]
    lib.ref_center_of_mass.argtypes = [dp, C.c_int64, dp]
    lib.ref_radius_of_gyration.restype = C.c_double
    lib.ref_radius_of_gyration.argtypes = [dp, C.c_int64]
    lib.ref_max_threads.restype = C.c_int
    lib.ref_set_threads.argtypes = [C.c_int]
    _REF_LIBS[key] = lib
    return lib


class _CpuSystemBase:
    """Common numpy-facing surface of RefSystem / OracleSystem."""

    _prefix = ""

    def _fn(self, name):
        return getattr(self.lib, self._prefix + name)

    def update_neighbor_list(self):
        self._fn("update_neighbor_list")(self.h)

    def do_step(self, dt: float):
        self._fn("do_step")(self.h, float(dt))

    def run(self, dt: float, n_steps: int, rebuild_every: int = 0):
        self._fn("run")(self.h, float(dt), int(n_steps), int(rebuild_every))

    def facet_update_positions(self, dt: float):
        self._fn("facet_update_positions")(self.h, float(dt))

    def get(self, name: str) -> np.ndarray:
        out = np.empty((self.n, 3), dtype=np.float64)
        self._fn("get_field")(self.h, FIELDS[name], _dp(out))
        return out

    def n_pairs(self) -> int:
        return int(self._fn("n_pairs")(self.h))

    def rows(self):
        p = self.n_pairs()
        row_ptr = np.empty(self.n + 1, dtype=np.int64)
        cols = np.empty(max(p, 1), dtype=np.int64)
        self._fn("get_rows")(self.h, _ip(row_ptr), _ip(cols))
        return row_ptr, cols[:p]

    def contact_springs(self, i: int, j: int) -> np.ndarray:
        out = np.zeros(9, dtype=np.float64)
        self._fn("get_contact_springs")(self.h, int(i), int(j), _dp(out))
        return out

    def bond(self, i: int, j: int):
        conn = np.zeros((5, 3), dtype=np.float64)
        length = np.zeros(5, dtype=np.float64)
        ok = self._fn("get_bond")(self.h, int(i), int(j), _dp(conn), _dp(length))
        return (conn, length) if ok else None

    def facet_vertices(self) -> np.ndarray:
        out = np.zeros((3, 3), dtype=np.float64)
        self._fn("get_facet_vertices")(self.h, _dp(out))
        return out

    def close(self):
        if getattr(self, "h", None):
            self._fn("destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class RefSystem(_CpuSystemBase):
    """One instance of the reference's granular_system[_neighbor_list] (oracle/ref_harness.cpp)."""

    _prefix = "ref_"

    def __init__(self, cfg, x0, v0, theta0, omega0, timing: bool = False):
        self.lib = load_ref(timing)
        self.n = int(cfg.n)
        self.cfg_struct = to_struct(cfg)
        x0, v0, theta0, omega0 = (_c3(a, self.n) for a in (x0, v0, theta0, omega0))
        self.h = self.lib.ref_create(C.byref(self.cfg_struct), _dp(x0), _dp(v0), _dp(theta0), _dp(omega0))
        if not self.h:
            raise RuntimeError("ref_create failed (unsupported model combination?)")


def ref_diagnostics(x, v, omega, mass, inertia, timing: bool = False) -> dict:
    """The reference's own diagnostics (test/compute_energy.cpp:7-58, test/mass_distribution.cpp:7-26)."""
    lib = load_ref(timing)
    n = len(x)
    x, v, omega = (_c3(a, n) for a in (x, v, omega))
    com = np.zeros(3)
    lib.ref_center_of_mass(_dp(x), n, _dp(com))
    return {
        "ke": lib.ref_total_kinetic_energy(_dp(v), _dp(omega), n, mass, inertia),
        "p": lib.ref_linear_momentum(_dp(v), n, mass),
        "l": lib.ref_angular_momentum(_dp(x), _dp(v), _dp(omega), n, mass, inertia),
        "com": com,
        "rg": lib.ref_radius_of_gyration(_dp(x), n),
    }


# ----------------------------------------------------------------------------------------------
# oracle/libdem_oracle.so : the CPU restatement
# ----------------------------------------------------------------------------------------------
_ORACLE_LIB = None


def oracle_lib_path() -> str:
    return os.path.join(_HERE, "libdem_oracle.so")


def build_oracle(force: bool = False) -> str:
    """Compile oracle/dem_oracle.cpp (plain g++; needs nothing from /root/reference)."""
    path = oracle_lib_path()
    src = os.path.join(_HERE, "dem_oracle.cpp")
    deps = [src, os.path.join(_HERE, "dem_oracle.h"), os.path.join(_HERE, "dem_config.h")]
    if force or not os.path.exists(path) or any(os.path.getmtime(d) > os.path.getmtime(path) for d in deps):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return path


def load_oracle():
    global _ORACLE_LIB
    if _ORACLE_LIB is not None:
        return _ORACLE_LIB
    lib = C.CDLL(build_oracle())
    dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_void_p
    lib.orc_create.restype = vp
    lib.orc_create.argtypes = [C.POINTER(DemConfigStruct), dp, dp, dp, dp]
    lib.orc_destroy.argtypes = [vp]
    lib.orc_update_neighbor_list.argtypes = [vp]
    lib.orc_do_step.argtypes = [vp, C.c_double]
    lib.orc_run.argtypes = [vp, C.c_double, C.c_int64, C.c_int64]
    lib.orc_facet_update_positions.argtypes = [vp, C.c_double]
    lib.orc_get_field.argtypes = [vp, C.c_int, dp]
    lib.orc_n_pairs.restype = C.c_int64
    lib.orc_n_pairs.argtypes = [vp]
    lib.orc_get_rows.argtypes = [vp, ip, ip]
    lib.orc_get_contact_springs.argtypes = [vp, C.c_int64, C.c_int64, dp]
    lib.orc_get_bond.restype = C.c_int
    lib.orc_get_bond.argtypes = [vp, C.c_int64, C.c_int64, dp, dp]
    lib.orc_get_facet_vertices.argtypes = [vp, dp]
    lib.orc_n_bonds.restype = C.c_int64
    lib.orc_n_bonds.argtypes = [vp]
    lib.orc_n_contacts.restype = C.c_int64
    lib.orc_n_contacts.argtypes = [vp]
    lib.orc_set_brute_force_neighbors.argtypes = [vp, C.c_int]
    lib.orc_unit_vectors.argtypes = [C.c_uint64, C.c_int64, dp]
    lib.orc_diagnostics.argtypes = [dp, dp, dp, C.c_int64, C.c_double, C.c_double, dp]
    lib.orc_max_threads.restype = C.c_int
    lib.orc_set_threads.argtypes = [C.c_int]
    _ORACLE_LIB = lib
    return lib


class OracleSystem(_CpuSystemBase):
    """CPU restatement with exposed sparse state (oracle/dem_oracle.cpp)."""

    _prefix = "orc_"

    def __init__(self, cfg, x0, v0, theta0, omega0):
        self.lib = load_oracle()
        self.n = int(cfg.n)
        self.cfg_struct = to_struct(cfg)
        x0, v0, theta0, omega0 = (_c3(a, self.n) for a in (x0, v0, theta0, omega0))
        self.h = self.lib.orc_create(C.byref(self.cfg_struct), _dp(x0), _dp(v0), _dp(theta0), _dp(omega0))
        if not self.h:
            raise RuntimeError("orc_create failed")

    def n_bonds(self) -> int:
        return int(self.lib.orc_n_bonds(self.h))

    def n_contacts(self) -> int:
        return int(self.lib.orc_n_contacts(self.h))

    def set_brute_force_neighbors(self, on: bool):
        self.lib.orc_set_brute_force_neighbors(self.h, int(bool(on)))


def oracle_unit_vectors(seed: int, count: int) -> np.ndarray:
    """``count`` unit vectors from mt19937_64(seed) + uniform(-1, 1), as test/sintered.cpp:29-39."""
    out = np.empty((count, 3), dtype=np.float64)
    load_oracle().orc_unit_vectors(int(seed), int(count), _dp(out))
    return out


def oracle_diagnostics(x, v, omega, mass, inertia) -> dict:
    """Restatement of test/compute_energy.cpp:7-58 and test/mass_distribution.cpp:7-26 (serial, index order)."""
    n = len(x)
    x, v, omega = (_c3(a, n) for a in (x, v, omega))
    out = np.zeros(8, dtype=np.float64)
    load_oracle().orc_diagnostics(_dp(x), _dp(v), _dp(omega), n, float(mass), float(inertia), _dp(out))
    return {"ke": out[0], "p": out[1], "l": out[2], "com": out[3:6].copy(), "rg": out[6]}
