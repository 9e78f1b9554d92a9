"""Python host-side mirror of libgran's system interface on top of the C ABI (include/lgb200/lgb200.h).

``GranularSystem`` follows the reference's driver-facing surface
(LG/granular_system/granular_system.h:53-81, granular_system_neighbor_list.h:54-82,
LT/rotational_system/rotational_system.h:83-121): construct with x0, v0, theta0, omega0, then
``do_step(dt)``, ``update_neighbor_list()``, ``get_x()`` ... ``get_alpha()``.  All computation happens in
the CUDA library; this module only marshals numpy arrays through ctypes.  If ``liblgb200.so`` is missing
or no GPU is present the calls raise -- there is no fallback path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Callable, Optional, Sequence

import numpy as np

from .build import LIB_PATH

LG_OK = 0
LG_EINVAL, LG_ENODEVICE, LG_ECUDA, LG_ENOMEM, LG_ESIZE, LG_ECAPACITY = -1, -2, -3, -4, -5, -6
FIELDS = {"x": 0, "v": 1, "a": 2, "theta": 3, "omega": 4, "alpha": 5}

#: every symbol declared in include/lgb200/lgb200.h (checked by tests/test_abi.py)
ABI_SYMBOLS = (
    "lg_create", "lg_destroy", "lg_last_error", "lg_abi_version", "lg_set_state", "lg_get_field", "lg_device_field",
    "lg_set_contact", "lg_set_hamaker", "lg_build_sinter_bonds", "lg_set_bonds", "lg_set_max_contacts", "lg_add_facet",
    "lg_facet_set_velocity", "lg_facet_update_positions", "lg_facet_get_vertices", "lg_enable_neighbor_list",
    "lg_update_neighbor_list", "lg_do_step", "lg_run", "lg_reset_acceleration_buffers", "lg_synchronize",
    "lg_get_neighbor_count", "lg_get_neighbor_rows", "lg_get_contact_springs", "lg_get_bond", "lg_get_counters",
    "lg_diagnostics", "lg_set_kernel_variant", "lg_get_stream", "lg_set_profiling", "lg_get_profile",
    "lg_bond_table_create", "lg_bond_table_count", "lg_bond_table_records", "lg_bond_table_destroy", "lg_set_bond_table",
    "lg_facet_set_vertices", "lg_set_math_mode", "lg_last_step_kernel",
    "lg_gen_fractal_aggregate", "lg_gen_random_packing",
    # SURVEY.md section 8f variants
    "lg_set_hamaker_variant", "lg_set_integrator", "lg_set_fixed_particle", "lg_set_central_attraction",
    "lg_alt_bond_table_create", "lg_set_alt_bond_table", "lg_get_alt_bond", "lg_set_reorder", "lg_set_autotune", "lg_get_rows_shape", "lg_bond_table_create_device", "lg_set_rows_shape",
    "lg_rows_shape_rule",
    "lg_dumper_create", "lg_dump_particles", "lg_dump_necks", "lg_dump_spring_connectors", "lg_dump_particles_csv",
    "lg_dumper_wait", "lg_dumper_destroy", "lg_write_particles", "lg_write_necks", "lg_write_spring_connectors",
    # slab decomposition support
    "lg_flush", "lg_first_half_kick", "lg_import_state_device", "lg_set_owned_device", "lg_export_slots",
    "lg_import_slots_device", "lg_export_bonds", "lg_import_bonds_device", "lg_halo_pack", "lg_halo_unpack",
    "lg_halo_create", "lg_halo_set_lists", "lg_halo_export", "lg_halo_connect", "lg_halo_exchange", "lg_run_slab",
    "lg_halo_disconnect", "lg_halo_destroy", "lg_export_wall", "lg_import_wall_device", "lg_diagnostics_partial",
    # no counterpart in the reference: allocation-free diagnostics sums (static device scratch)
    "lg_diag_sums_static", "lg_diag_sums_static_max_particles",
)


class SizeMismatchException(Exception):
    """LT/exception/exception.h:11-24."""


class LgError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"lgb200 error {code}: {msg}")
        self.code = code


class ContactParamsStruct(C.Structure):
    _fields_ = [(name, C.c_double) for name in (
        "k", "gamma_n", "k_t", "gamma_t", "mu_s", "phi_d", "k_r", "gamma_r", "mu_r", "phi_r", "k_o", "gamma_o", "mu_o",
        "phi_o", "dt")]

    @classmethod
    def from_seq(cls, seq: Sequence[float]) -> "ContactParamsStruct":
        if len(seq) != 15:
            raise ValueError("contact parameter tuple needs 15 entries")
        return cls(*[float(v) for v in seq])


UNIT_VECTOR_FN = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_double))

_LIB = None


def load_library(path: Optional[str] = None):
    """dlopen the C-ABI library and declare its prototypes. Raises if it has not been built."""
    global _LIB
    if _LIB is not None and path is None:
        return _LIB
    # LGB200_LIB selects an experimental build of the same ABI (kernel tuning sweeps); never a CPU path
    path = path or os.environ.get("LGB200_LIB") or LIB_PATH
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(libgran_b200 has no CPU fallback)")
    lib = C.CDLL(path)
    dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int64), C.c_void_p
    cpp = C.POINTER(ContactParamsStruct)
    lib.lg_last_error.restype = C.c_char_p
    lib.lg_abi_version.restype = C.c_int
    lib.lg_create.argtypes = [C.c_int64, C.c_double, C.c_double, C.c_double, C.c_int, C.POINTER(vp)]
    lib.lg_destroy.argtypes = [vp]
    lib.lg_set_state.argtypes = [vp, C.c_int64, dp, dp, dp, dp]
    lib.lg_get_field.argtypes = [vp, C.c_int, dp]
    lib.lg_device_field.argtypes = [vp, C.c_int, C.POINTER(vp)]
    lib.lg_set_contact.argtypes = [vp, cpp]
    lib.lg_set_hamaker.argtypes = [vp, C.c_double, C.c_double]
    lib.lg_build_sinter_bonds.argtypes = [vp, cpp, C.c_double, C.c_double, C.c_double, UNIT_VECTOR_FN, vp, ip]
    lib.lg_set_bonds.argtypes = [vp, cpp, C.c_double, C.c_double, C.c_int64, ip, dp, dp]
    lib.lg_set_max_contacts.argtypes = [vp, C.c_int]
    lib.lg_add_facet.argtypes = [vp, dp, dp, cpp, C.c_int, C.c_double, C.c_double, C.POINTER(C.c_int)]
    lib.lg_facet_set_velocity.argtypes = [vp, C.c_int, dp]
    lib.lg_facet_update_positions.argtypes = [vp, C.c_int, C.c_double]
    lib.lg_facet_get_vertices.argtypes = [vp, C.c_int, dp]
    lib.lg_enable_neighbor_list.argtypes = [vp, C.c_double]
    lib.lg_update_neighbor_list.argtypes = [vp]
    lib.lg_do_step.argtypes = [vp, C.c_double]
    lib.lg_run.argtypes = [vp, C.c_double, C.c_int64, C.c_int64, C.c_int]
    lib.lg_reset_acceleration_buffers.argtypes = [vp]
    lib.lg_synchronize.argtypes = [vp]
    lib.lg_get_neighbor_count.argtypes = [vp, ip]
    lib.lg_get_neighbor_rows.argtypes = [vp, ip, ip]
    lib.lg_get_contact_springs.argtypes = [vp, C.c_int64, C.c_int64, dp]
    lib.lg_get_bond.argtypes = [vp, C.c_int64, C.c_int64, C.POINTER(C.c_int), dp, dp]
    lib.lg_get_counters.argtypes = [vp, ip]
    lib.lg_diagnostics.argtypes = [vp, dp]
    lib.lg_set_kernel_variant.argtypes = [vp, C.c_int]
    lib.lg_get_stream.argtypes = [vp, C.POINTER(vp)]
    lib.lg_set_profiling.argtypes = [vp, C.c_int]
    lib.lg_get_profile.argtypes = [vp, dp, ip]
    lib.lg_bond_table_create.argtypes = [C.c_int64, vp, C.c_double, C.c_double, UNIT_VECTOR_FN, vp, C.POINTER(vp)]
    lib.lg_bond_table_count.argtypes = [vp, ip]
    lib.lg_bond_table_records.argtypes = [vp, vp, vp, vp]
    lib.lg_bond_table_destroy.argtypes = [vp]
    lib.lg_set_bond_table.argtypes = [vp, cpp, C.c_double, C.c_double, vp]
    lib.lg_facet_set_vertices.argtypes = [vp, C.c_int, dp]
    lib.lg_set_math_mode.argtypes = [vp, C.c_int]
    lib.lg_last_step_kernel.argtypes = [vp, C.c_char_p, C.c_int]
    lib.lg_gen_fractal_aggregate.argtypes = [C.c_int64, C.c_uint64, C.c_double, dp]
    lib.lg_gen_random_packing.argtypes = [C.c_int64, C.c_double, C.c_double, C.c_double, C.c_uint64, C.c_double, dp, dp]
    lib.lg_set_hamaker_variant.argtypes = [vp, C.c_int, C.c_double]
    lib.lg_set_integrator.argtypes = [vp, C.c_int]
    lib.lg_set_fixed_particle.argtypes = [vp, C.c_int64]
    lib.lg_set_central_attraction.argtypes = [vp, C.c_double, dp]
    lib.lg_alt_bond_table_create.argtypes = [C.c_int64, vp, C.c_double, C.c_double, C.POINTER(vp)]
    lib.lg_set_alt_bond_table.argtypes = [vp, cpp, dp, vp]
    lib.lg_get_alt_bond.argtypes = [vp, C.c_int64, C.c_int64, C.POINTER(C.c_int), dp]
    lib.lg_set_reorder.argtypes = [vp, C.c_int]
    lib.lg_set_autotune.argtypes = [vp, C.c_int]
    lib.lg_bond_table_create_device.argtypes = [C.c_int64, vp, C.c_double, C.c_double, C.c_int64, vp, C.c_int, C.POINTER(vp)]
    lib.lg_get_rows_shape.argtypes = [vp, C.POINTER(C.c_int)]
    lib.lg_set_rows_shape.argtypes = [vp, C.c_int]
    lib.lg_rows_shape_rule.argtypes = [C.c_double, C.c_int64, C.c_int]
    lib.lg_dumper_create.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    for name in ("lg_dump_particles", "lg_dump_necks", "lg_dump_spring_connectors"):
        getattr(lib, name).argtypes = [vp, vp]
    lib.lg_dump_particles_csv.argtypes = [vp, vp, C.c_int64]
    lib.lg_dumper_wait.argtypes = [vp]
    lib.lg_dumper_destroy.argtypes = [vp]
    lib.lg_write_particles.argtypes = [C.c_char_p, C.c_int64, dp, dp, C.c_double]
    lib.lg_write_necks.argtypes = [C.c_char_p, C.c_int64, dp, C.c_double, C.c_int64, ip]
    lib.lg_write_spring_connectors.argtypes = [C.c_char_p, C.c_double, C.c_int64, dp]
    # slab decomposition support: device pointers travel as plain integers
    lib.lg_flush.argtypes = [vp]
    lib.lg_first_half_kick.argtypes = [vp, C.c_double]
    lib.lg_import_state_device.argtypes = [vp, vp, vp, vp, vp, vp, vp, C.c_int]
    lib.lg_set_owned_device.argtypes = [vp, vp]
    lib.lg_export_slots.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(vp), C.POINTER(vp)]
    lib.lg_import_slots_device.argtypes = [vp, vp, vp]
    lib.lg_export_bonds.argtypes = [vp, ip, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    lib.lg_import_bonds_device.argtypes = [vp, cpp, C.c_double, C.c_double, C.c_int64, vp, vp, vp, vp, vp]
    lib.lg_halo_pack.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp]
    lib.lg_halo_unpack.argtypes = [vp, C.c_int, vp, C.c_int, vp, vp]
    lib.lg_halo_create.argtypes = [vp, C.c_int, C.POINTER(C.c_int64), C.POINTER(vp)]
    lib.lg_halo_set_lists.argtypes = [vp, C.c_int, C.c_int, vp, C.c_int, vp, C.c_int, vp, C.c_int, vp]
    lib.lg_halo_export.argtypes = [vp, C.c_int, C.c_char_p, C.POINTER(C.c_int64)]
    lib.lg_halo_connect.argtypes = [vp, C.c_int, C.c_char_p, C.POINTER(C.c_int64)]
    lib.lg_halo_exchange.argtypes = [vp, vp]
    lib.lg_run_slab.argtypes = [vp, vp, C.c_double, C.c_int64, C.c_int]
    lib.lg_export_wall.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(vp)]
    lib.lg_import_wall_device.argtypes = [vp, C.c_int, vp, vp]
    lib.lg_diagnostics_partial.argtypes = [vp, C.c_int, vp, C.POINTER(vp)]
    lib.lg_diag_sums_static.argtypes = [C.c_int64, vp, vp, vp, C.c_double, C.c_double, C.c_int, vp, vp, vp]
    lib.lg_diag_sums_static_max_particles.argtypes = []
    lib.lg_halo_disconnect.argtypes = [vp]
    lib.lg_halo_destroy.argtypes = [vp]
    for name in ABI_SYMBOLS:
        fn = getattr(lib, name)
        if name not in ("lg_last_error",):
            fn.restype = C.c_int if name != "lg_last_error" else C.c_char_p
    lib.lg_last_error.restype = C.c_char_p
    if path == (os.environ.get("LGB200_LIB") or LIB_PATH):
        _LIB = lib
    return lib


def _dp(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a: np.ndarray):
    return a.ctypes.data_as(C.POINTER(C.c_int64))


def _c3(a, n=None) -> np.ndarray:
    out = np.ascontiguousarray(np.asarray(a, dtype=np.float64))
    if out.ndim != 2 or out.shape[1] != 3:
        out = out.reshape(-1, 3)
    return out


class GranularSystem:
    """One device-resident granular system (handle of the C ABI)."""

    def __init__(self, x0, v0, theta0, omega0, r_part: float, mass: float, inertia: float, device: int = -1,
                 r_verlet: Optional[float] = None, reorder: bool = False):
        self.lib = load_library()
        x0, v0, theta0, omega0 = (_c3(a) for a in (x0, v0, theta0, omega0))
        # rotational_system.h:69-70
        if not (len(x0) == len(v0) == len(theta0) == len(omega0)):
            raise SizeMismatchException("size mismatch in arguments passed to GranularSystem(x0, v0, theta0, omega0)")
        self.n = len(x0)
        self.h = C.c_void_p()
        self._facets = 0
        self._cb_keepalive = []
        self._check(self.lib.lg_create(self.n, float(r_part), float(mass), float(inertia), int(device), C.byref(self.h)))
        if reorder:  # internal spatial renumbering (lg_set_reorder): invisible here, every call speaks the caller's order
            self._check(self.lib.lg_set_reorder(self.h, 1))
        self._check(self.lib.lg_set_state(self.h, self.n, _dp(x0), _dp(v0), _dp(theta0), _dp(omega0)))
        if r_verlet is not None:
            self.enable_neighbor_list(r_verlet)

    # -- plumbing ------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != LG_OK:
            msg = self.lib.lg_last_error().decode(errors="replace")
            if rc == LG_ESIZE:
                raise SizeMismatchException(msg)
            raise LgError(rc, msg)

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.lib.lg_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- model installation ----------------------------------------------------------------------
    def set_state(self, x, v, theta, omega):
        x, v, theta, omega = (_c3(a) for a in (x, v, theta, omega))
        if not (len(x) == len(v) == len(theta) == len(omega)):
            raise SizeMismatchException("size mismatch in arguments passed to set_state")
        self._check(self.lib.lg_set_state(self.h, len(x), _dp(x), _dp(v), _dp(theta), _dp(omega)))

    def set_contact(self, params: Sequence[float]):
        p = ContactParamsStruct.from_seq(params)
        self._check(self.lib.lg_set_contact(self.h, C.byref(p)))

    def set_hamaker(self, A: float, h0: float):
        self._check(self.lib.lg_set_hamaker(self.h, float(A), float(h0)))

    def set_hamaker_variant(self, kind: int, r_cutoff: float = 0.0):
        """1 = limited_range_hamaker_functor (zero beyond r_cutoff), 2 = simplified_hamaker_functor."""
        self._check(self.lib.lg_set_hamaker_variant(self.h, int(kind), float(r_cutoff)))

    def set_integrator(self, integrator: int):
        """0 = rotational_velocity_verlet_half, 1 = rotational_forward_euler."""
        self._check(self.lib.lg_set_integrator(self.h, int(integrator)))

    def set_fixed_particle(self, index: int):
        self._check(self.lib.lg_set_fixed_particle(self.h, int(index)))

    def set_central_attraction(self, k_a: float, x0):
        c = np.ascontiguousarray(np.asarray(x0, dtype=np.float64).reshape(3))
        self._check(self.lib.lg_set_central_attraction(self.h, float(k_a), _dp(c)))

    def build_alt_sinter_bonds(self, contact: Sequence[float], bond_params: Sequence[float], critical_separation: float,
                               x0, r_part: float) -> int:
        """alt_sinter_functor's constructor: bond scan over x0 (cycle-free), embedded contact model, and the bond's own
        {k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o, dt}."""
        p = ContactParamsStruct.from_seq(contact)
        bp = np.ascontiguousarray(np.asarray(bond_params, dtype=np.float64).reshape(9))
        x = _c3(x0)
        tbl = C.c_void_p()
        self._check(self.lib.lg_alt_bond_table_create(len(x), x.ctypes.data, float(r_part), float(critical_separation), C.byref(tbl)))
        try:
            nb = C.c_int64(0)
            self.lib.lg_bond_table_count(tbl, C.byref(nb))
            self._check(self.lib.lg_set_alt_bond_table(self.h, C.byref(p), _dp(bp), tbl))
        finally:
            self.lib.lg_bond_table_destroy(tbl)
        return int(nb.value)

    def alt_bond(self, i: int, j: int):
        out, flag = np.zeros(9), C.c_int(0)
        self._check(self.lib.lg_get_alt_bond(self.h, int(i), int(j), C.byref(flag), _dp(out)))
        return out if flag.value else None

    def set_autotune(self, on: bool = True):
        self._check(self.lib.lg_set_autotune(self.h, int(bool(on))))

    def rows_shape(self) -> int:
        v = C.c_int(-1)
        self._check(self.lib.lg_get_rows_shape(self.h, C.byref(v)))
        return int(v.value)

    def set_max_contacts(self, m: int):
        self._check(self.lib.lg_set_max_contacts(self.h, int(m)))

    def build_sinter_bonds(self, contact: Sequence[float], k_d: float, gamma_d: float, critical_separation: float,
                           unit_vectors) -> int:
        """``unit_vectors`` is either a callable returning a 3-vector (the reference's
        random_unit_vector_generator) or an (m, 3) array consumed in order."""
        p = ContactParamsStruct.from_seq(contact)
        if callable(unit_vectors):
            def cb(_user, out):
                vec = unit_vectors()
                out[0], out[1], out[2] = float(vec[0]), float(vec[1]), float(vec[2])
        else:
            arr = _c3(unit_vectors)
            state = {"k": 0}

            def cb(_user, out):
                k = state["k"]
                if k >= len(arr):  # an exception cannot cross the C frame: hand out a valid vector, fail after the call
                    state["short"] = True
                    out[0], out[1], out[2] = 1.0, 0.0, 0.0
                    return
                out[0], out[1], out[2] = arr[k, 0], arr[k, 1], arr[k, 2]
                state["k"] = k + 1
        cfn = UNIT_VECTOR_FN(cb)
        self._cb_keepalive.append(cfn)
        nb = C.c_int64(0)
        self._check(self.lib.lg_build_sinter_bonds(self.h, C.byref(p), float(k_d), float(gamma_d),
                                                   float(critical_separation), cfn, None, C.byref(nb)))
        if not callable(unit_vectors) and state.get("short"):
            raise IndexError(f"unit vector array exhausted: {len(arr)} vectors for {int(nb.value)} bonds")
        return int(nb.value)

    def set_bonds(self, contact: Sequence[float], k_d: float, gamma_d: float, ij, connectors, rest_lengths):
        p = ContactParamsStruct.from_seq(contact)
        ij = np.ascontiguousarray(np.asarray(ij, dtype=np.int64).reshape(-1, 2))
        conn = np.ascontiguousarray(np.asarray(connectors, dtype=np.float64).reshape(len(ij), 15))
        rl = np.ascontiguousarray(np.asarray(rest_lengths, dtype=np.float64).reshape(len(ij), 5))
        self._check(self.lib.lg_set_bonds(self.h, C.byref(p), float(k_d), float(gamma_d), len(ij), _ip(ij), _dp(conn), _dp(rl)))

    def add_facet(self, vertices, v_facet, surface_contact: Optional[Sequence[float]] = None,
                  hamaker: Optional[Sequence[float]] = None) -> int:
        verts = np.ascontiguousarray(np.asarray(vertices, dtype=np.float64).reshape(9))
        vf = np.ascontiguousarray(np.asarray(v_facet, dtype=np.float64).reshape(3))
        p = ContactParamsStruct.from_seq(surface_contact) if surface_contact is not None else None
        A, h0 = (float(hamaker[0]), float(hamaker[1])) if hamaker is not None else (0.0, 0.0)
        fid = C.c_int(-1)
        self._check(self.lib.lg_add_facet(self.h, _dp(verts), _dp(vf), C.byref(p) if p is not None else None,
                                          int(hamaker is not None), A, h0, C.byref(fid)))
        self._facets += 1
        return int(fid.value)

    def facet_set_velocity(self, facet_id: int, v):
        vf = np.ascontiguousarray(np.asarray(v, dtype=np.float64).reshape(3))
        self._check(self.lib.lg_facet_set_velocity(self.h, int(facet_id), _dp(vf)))

    def facet_update_positions(self, dt: float, facet_id: int = 0):
        self._check(self.lib.lg_facet_update_positions(self.h, int(facet_id), float(dt)))

    def facet_vertices(self, facet_id: int = 0) -> np.ndarray:
        out = np.zeros((3, 3))
        self._check(self.lib.lg_facet_get_vertices(self.h, int(facet_id), _dp(out)))
        return out

    def enable_neighbor_list(self, r_verlet: float):
        self._check(self.lib.lg_enable_neighbor_list(self.h, float(r_verlet)))

    def set_kernel_variant(self, variant: int):
        self._check(self.lib.lg_set_kernel_variant(self.h, int(variant)))

    # -- the reference's driver-facing calls --------------------------------------------------------
    def set_math_mode(self, mode: int):
        """0 = bit-exact (default), 1 = relaxed arithmetic (see lg_set_math_mode)."""
        self._check(self.lib.lg_set_math_mode(self.h, int(mode)))

    def update_neighbor_list(self):
        self._check(self.lib.lg_update_neighbor_list(self.h))

    def do_step(self, dt: float):
        self._check(self.lib.lg_do_step(self.h, float(dt)))

    def run(self, dt: float, n_steps: int, rebuild_every: int = 0, move_facets: bool = False):
        self._check(self.lib.lg_run(self.h, float(dt), int(n_steps), int(rebuild_every), int(move_facets)))

    def reset_acceleration_buffers(self):
        self._check(self.lib.lg_reset_acceleration_buffers(self.h))

    def synchronize(self):
        self._check(self.lib.lg_synchronize(self.h))

    def get(self, name: str, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty((self.n, 3), dtype=np.float64)
        self._check(self.lib.lg_get_field(self.h, FIELDS[name], _dp(out)))
        return out

    def get_x(self): return self.get("x")
    def get_v(self): return self.get("v")
    def get_a(self): return self.get("a")
    def get_theta(self): return self.get("theta")
    def get_omega(self): return self.get("omega")
    def get_alpha(self): return self.get("alpha")

    def stream_handle(self) -> int:
        """cudaStream_t of the system, e.g. for torch.cuda.ExternalStream."""
        p = C.c_void_p()
        self._check(self.lib.lg_get_stream(self.h, C.byref(p)))
        return int(p.value or 0)

    def set_profiling(self, on: bool):
        self._check(self.lib.lg_set_profiling(self.h, int(bool(on))))

    def get_profile(self):
        ms, cnt = C.c_double(0.0), C.c_int64(0)
        self._check(self.lib.lg_get_profile(self.h, C.byref(ms), C.byref(cnt)))
        return float(ms.value), int(cnt.value)

    def last_step_kernel(self) -> str:
        buf = C.create_string_buffer(64)
        self._check(self.lib.lg_last_step_kernel(self.h, buf, 64))
        return buf.value.decode()

    def device_ptr(self, name: str) -> int:
        p = C.c_void_p()
        self._check(self.lib.lg_device_field(self.h, FIELDS[name], C.byref(p)))
        return int(p.value)

    # -- introspection ----------------------------------------------------------------------------
    def n_pairs(self) -> int:
        v = C.c_int64(0)
        self._check(self.lib.lg_get_neighbor_count(self.h, C.byref(v)))
        return int(v.value)

    def rows(self):
        p = self.n_pairs()
        row_ptr = np.empty(self.n + 1, dtype=np.int64)
        cols = np.empty(max(p, 1), dtype=np.int64)
        self._check(self.lib.lg_get_neighbor_rows(self.h, _ip(row_ptr), _ip(cols)))
        return row_ptr, cols[:p]

    def contact_springs(self, i: int, j: int) -> np.ndarray:
        out = np.zeros(9)
        self._check(self.lib.lg_get_contact_springs(self.h, int(i), int(j), _dp(out)))
        return out

    def bond(self, i: int, j: int):
        conn, length, flag = np.zeros((5, 3)), np.zeros(5), C.c_int(0)
        self._check(self.lib.lg_get_bond(self.h, int(i), int(j), C.byref(flag), _dp(conn), _dp(length)))
        return (conn, length) if flag.value else None

    def counters(self) -> dict:
        out = np.zeros(8, dtype=np.int64)
        self._check(self.lib.lg_get_counters(self.h, _ip(out)))
        keys = ("pairs", "contacts", "bonds", "steps", "rebuilds", "launches", "cells", "_")
        return dict(zip(keys, (int(v) for v in out)))

    def diagnostics(self) -> dict:
        out = np.zeros(9)
        self._check(self.lib.lg_diagnostics(self.h, _dp(out)))
        return {"ke": out[0], "ke_trans": out[1], "ke_rot": out[2], "p": out[3], "l": out[4], "com": out[5:8].copy(),
                "rg": out[8]}


class Dumper:
    """The dump pipeline of a system (lg_dumper_*): VTK files as the reference's writer.cpp writes them, produced off the
    time loop (device snapshot -> copy stream -> background writer thread)."""

    def __init__(self, system: GranularSystem, directory: str):
        self.sys = system
        self.lib = system.lib
        self.h = C.c_void_p()
        system._check(self.lib.lg_dumper_create(system.h, os.fsencode(directory), C.byref(self.h)))

    def particles(self):
        self.sys._check(self.lib.lg_dump_particles(self.sys.h, self.h))

    def necks(self):
        self.sys._check(self.lib.lg_dump_necks(self.sys.h, self.h))

    def spring_connectors(self):
        self.sys._check(self.lib.lg_dump_spring_connectors(self.sys.h, self.h))

    def particles_csv(self, count: int):
        self.sys._check(self.lib.lg_dump_particles_csv(self.sys.h, self.h, int(count)))

    def wait(self):
        self.sys._check(self.lib.lg_dumper_wait(self.h))

    def close(self):
        if self.h:
            self.lib.lg_dumper_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def from_config(cfg, x0, v0, theta0, omega0, unit_vectors=None, device: int = -1, variant: Optional[int] = None,
                math_mode: int = 0, bonds=None, reorder: bool = False) -> GranularSystem:
    """Build a GranularSystem from a ``scenarios.DemConfig`` (same meaning as oracle/dem_config.h)."""
    from .scenarios import MODEL_ALT_SINTER, MODEL_CONTACT, MODEL_FACET, MODEL_HAMAKER, MODEL_SINTER

    sysm = GranularSystem(x0, v0, theta0, omega0, cfg.r_part, cfg.mass, cfg.inertia, device=device,
                          r_verlet=cfg.r_verlet if cfg.use_neighbor_list else None, reorder=reorder)
    if variant is not None:
        sysm.set_kernel_variant(variant)
    if math_mode:
        sysm.set_math_mode(math_mode)
    if cfg.model_mask & MODEL_CONTACT:
        sysm.set_contact(cfg.contact)
    if cfg.model_mask & MODEL_SINTER:
        if bonds is not None:  # explicit tables (ij, connectors, rest lengths) instead of the constructor's scan
            sysm.set_bonds(cfg.contact, cfg.k_d, cfg.gamma_d, *bonds)
        else:
            if unit_vectors is None:
                raise ValueError("the sinter model needs unit_vectors (callable or array)")
            sysm.build_sinter_bonds(cfg.contact, cfg.k_d, cfg.gamma_d, cfg.critical_separation, unit_vectors)
    if cfg.model_mask & MODEL_ALT_SINTER:
        sysm.build_alt_sinter_bonds(cfg.contact, tuple(cfg.alt_bond) + (cfg.contact[14],), cfg.critical_separation, x0, cfg.r_part)
    if cfg.model_mask & MODEL_HAMAKER:
        sysm.set_hamaker(cfg.hamaker_A, cfg.hamaker_h0)
        if getattr(cfg, "hamaker_kind", 0):
            sysm.set_hamaker_variant(cfg.hamaker_kind, cfg.hamaker_cutoff)
    if getattr(cfg, "integrator", 0):
        sysm.set_integrator(cfg.integrator)
    if getattr(cfg, "fixed_particle", -1) >= 0:
        sysm.set_fixed_particle(cfg.fixed_particle)
    if cfg.model_mask & MODEL_FACET:
        sysm.add_facet(cfg.facet_vertices, cfg.facet_v, cfg.surf_contact, (cfg.surf_A, cfg.surf_h0))
    return sysm


def _table_records(lib, tbl):
    nb = C.c_int64(0)
    lib.lg_bond_table_count(tbl, C.byref(nb))
    m = 2 * nb.value
    ij = np.zeros((m, 2), dtype=np.int64)
    conn = np.zeros((m, 15), dtype=np.float64)
    ln = np.zeros((m, 5), dtype=np.float64)
    if m:
        lib.lg_bond_table_records(tbl, ij.ctypes.data, conn.ctypes.data, ln.ctypes.data)
    return ij, conn, ln


def device_bond_table(x, r_part: float, critical_separation: float, unit_vectors, device: int = -1):
    """The table of host_bond_table built on the GPU (lg_bond_table_create_device): cell-list candidates, bond k in
    lexicographic pair order takes unit_vectors[k].  Same records bit for bit."""
    lib = load_library()
    x = _c3(x)
    uv = _c3(unit_vectors)
    tbl = C.c_void_p()
    rc = lib.lg_bond_table_create_device(len(x), x.ctypes.data, float(r_part), float(critical_separation), len(uv), uv.ctypes.data,
                                         int(device), C.byref(tbl))
    if rc != LG_OK:
        raise LgError(rc, lib.lg_last_error().decode(errors="replace"))
    try:
        return _table_records(lib, tbl)
    finally:
        lib.lg_bond_table_destroy(tbl)


def host_bond_table(x, r_part: float, critical_separation: float, unit_vectors):
    """sinter_functor's constructor scan (LG/sinter_bridge/sinter_bridge.h:148-211) through lg_bond_table_create:
    pure host code, no GPU needed.  Returns (ij int64 [2b, 2], connectors [2b, 15], rest lengths [2b, 5]) --
    both orientations of every bond, in the layout lg_set_bonds takes."""
    lib = load_library()
    x = _c3(x)
    if callable(unit_vectors):
        def cb(_user, out):
            vec = unit_vectors()
            out[0], out[1], out[2] = float(vec[0]), float(vec[1]), float(vec[2])
    else:
        arr = _c3(unit_vectors)
        state = {"k": 0}

        def cb(_user, out):
            k = state["k"]
            out[0], out[1], out[2] = arr[k, 0], arr[k, 1], arr[k, 2]
            state["k"] = k + 1
    cfn = UNIT_VECTOR_FN(cb)
    tbl = C.c_void_p()
    rc = lib.lg_bond_table_create(len(x), x.ctypes.data, float(r_part), float(critical_separation), cfn, None, C.byref(tbl))
    if rc != LG_OK:
        raise LgError(rc, lib.lg_last_error().decode(errors="replace"))
    try:
        nb = C.c_int64(0)
        lib.lg_bond_table_count(tbl, C.byref(nb))
        m = 2 * nb.value
        ij = np.zeros((m, 2), dtype=np.int64)
        conn = np.zeros((m, 15), dtype=np.float64)
        ln = np.zeros((m, 5), dtype=np.float64)
        if m:
            lib.lg_bond_table_records(tbl, ij.ctypes.data, conn.ctypes.data, ln.ctypes.data)
    finally:
        lib.lg_bond_table_destroy(tbl)
    return ij, conn, ln
