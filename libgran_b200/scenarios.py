"""Initial conditions and parameter sets of the reference's test drivers and of BASELINE.json's configs.

Everything here is host-side numpy set-up code (what a libgran *driver program* does before the time
loop, e.g. test/hamaker_neighbor.cpp:28-79, test/sintered.cpp:52-104, test/surface.cpp:48-91,
main.cpp:56-93).  Nothing in this module computes forces or integrates.

Physical constants are the ones shared by all reference tests:
dt = 1e-13 s, rho = 1700, r = 1.4e-8 m, k = 1e4, gamma_n = 5e-9, ...
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass, field, replace
from typing import Tuple

import numpy as np

MODEL_CONTACT, MODEL_HAMAKER, MODEL_SINTER, MODEL_FACET, MODEL_ALT_SINTER = 1, 2, 4, 8, 16
HAMAKER_STANDARD, HAMAKER_LIMITED_RANGE, HAMAKER_SIMPLIFIED = 0, 1, 2
INTEGRATOR_VERLET_HALF, INTEGRATOR_FORWARD_EULER = 0, 1

R_PART = 1.4e-8
RHO = 1700.0
DT = 1e-13
MASS = 4.0 / 3.0 * math.pi * math.pow(R_PART, 3.0) * RHO
INERTIA = 2.0 / 5.0 * MASS * math.pow(R_PART, 2.0)


def contact_params(mu: float = 1.0, dt: float = DT) -> Tuple[float, ...]:
    """(k, gamma_n, k_t, gamma_t, mu_s, phi_d, k_r, gamma_r, mu_r, phi_r, k_o, gamma_o, mu_o, phi_o, dt)
    exactly as the tests pass them (test/hamaker_neighbor.cpp:41-49,87-89)."""
    k = 10000.0
    gamma_n = 5.0e-9
    phi = 1.0
    mu_o = 0.1
    gamma_t = 0.2 * gamma_n
    gamma_r = 0.05 * gamma_n
    gamma_o = 0.05 * gamma_n
    return (k, gamma_n, k, gamma_t, mu, phi, k, gamma_r, mu_o, phi, k, gamma_o, mu_o, phi, dt)


@dataclass
class DemConfig:
    """Python twin of ``struct dem_config`` (oracle/dem_config.h) and of the lg_* setter calls."""

    n: int
    model_mask: int
    use_neighbor_list: bool = False
    r_verlet: float = 5.0 * R_PART
    r_part: float = R_PART
    mass: float = MASS
    inertia: float = INERTIA
    contact: Tuple[float, ...] = field(default_factory=contact_params)
    hamaker_A: float = 1.0e-20
    hamaker_h0: float = 1.0e-9
    k_d: float = 10000.0
    gamma_d: float = 0.0
    critical_separation: float = 1.0e-9
    rng_seed: int = 0
    facet_vertices: np.ndarray = field(default_factory=lambda: np.zeros((3, 3)))
    facet_v: Tuple[float, float, float] = (0.0, 0.0, 0.0)
    surf_contact: Tuple[float, ...] = field(default_factory=lambda: contact_params(mu=0.5))
    surf_A: float = 0.0
    surf_h0: float = 1.0e-9
    dt: float = DT
    rebuild_every: int = 20
    # SURVEY.md section 8f variants
    alt_bond: Tuple[float, ...] = (0.0,) * 8   # alt_sinter_functor: k, gamma_n, k_t, gamma_t, k_r, gamma_r, k_o, gamma_o
    hamaker_kind: int = 0
    hamaker_cutoff: float = 0.0
    integrator: int = 0
    fixed_particle: int = -1

    def with_(self, **kw) -> "DemConfig":
        return replace(self, **kw)


@dataclass
class Scenario:
    name: str
    cfg: DemConfig
    x0: np.ndarray
    v0: np.ndarray
    theta0: np.ndarray
    omega0: np.ndarray
    #: optional: unit_vectors -> (ij, connectors, rest lengths) for lg_set_bonds, when the bridges are NOT what the sinter
    #: constructor would detect from x0 (multi_aggregate_in_contact: touching aggregates must stay unbonded)
    make_bonds: object = None

    @property
    def state(self):
        return self.x0, self.v0, self.theta0, self.omega0


def _zeros(n):
    return np.zeros((n, 3), dtype=np.float64)


# ---------------------------------------------------------------------------------------------
# the reference's own test set-ups
# ---------------------------------------------------------------------------------------------
def slab_and_projectile(models: int = MODEL_CONTACT | MODEL_HAMAKER, neighbor_list: bool = True) -> Scenario:
    """18-particle slab + one projectile (test/contact.cpp:56-77, test/hamaker.cpp, test/hamaker_neighbor.cpp:56-79)."""
    r = R_PART
    pts = []
    for i in range(3):
        y = -3.0 * r + float(i) * 2.0 * r
        for j in range(3):
            x = -3.0 * r + float(j) * 2.0 * r
            pts.append((x, y, 0.0))
            pts.append((x, y, -2.0 * r))
    pts.append((0.0, 0.0, 3.0 * r))
    x0 = np.array(pts, dtype=np.float64)
    n = len(x0)
    v0 = _zeros(n)
    v0[-1] = (1.0, 0.0, -10.0)
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=neighbor_list)
    return Scenario("slab_and_projectile", cfg, x0, v0, _zeros(n), _zeros(n))


def sintered_lattice(models: int = MODEL_SINTER, neighbor_list: bool = False) -> Scenario:
    """3x3x3 lattice, edge columns kicked (test/sintered.cpp:77-104)."""
    r = R_PART
    xs, vs = [], []
    for i in range(3):
        y = -1.0 * r + float(i) * 2.0 * r
        for j in range(3):
            x = -1.0 * r + float(j) * 2.0 * r
            xs += [(x, y, 0.0), (x, y, -2.0 * r), (x, y, -4.0 * r)]
            if j == 0:
                vel = (0.0, 0.0, 5.0)
            elif j == 2:
                vel = (0.0, 5.0, 0.0)
            else:
                vel = (0.0, 0.0, 0.0)
            vs += [vel] * 3
    x0 = np.array(xs, dtype=np.float64)
    v0 = np.array(vs, dtype=np.float64)
    n = len(x0)
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=neighbor_list, rng_seed=0)
    return Scenario("sintered_lattice", cfg, x0, v0, _zeros(n), _zeros(n))


def alt_sintered_lattice(neighbor_list: bool = False, hamaker: bool = False) -> Scenario:
    """3x3x3 lattice bonded by alt_sinter_functor (test/alt_sintered.cpp:43-113): dt = 5e-14, bond stiffness 1000 with
    critical damping, the embedded contact model with the common parameters."""
    base = sintered_lattice()
    dt = 5e-14
    k_bond = 1000.0
    gamma_n_bond = 2.0 * math.sqrt(2.0 * MASS * k_bond)
    alt = (k_bond, gamma_n_bond, k_bond, 0.2 * gamma_n_bond, k_bond, 0.05 * gamma_n_bond, k_bond, 0.05 * gamma_n_bond)
    models = MODEL_ALT_SINTER | (MODEL_HAMAKER if hamaker else 0)
    cfg = DemConfig(n=base.cfg.n, model_mask=models, use_neighbor_list=neighbor_list, contact=contact_params(dt=dt), dt=dt,
                    alt_bond=alt, critical_separation=1.0e-9)
    return Scenario("alt_sintered_lattice", cfg, base.x0, base.v0, base.theta0, base.omega0)


def surface_drop() -> Scenario:
    """One particle dropped on a triangular facet (test/surface.cpp:48-105); surface Hamaker A = 0."""
    r = R_PART
    a = 50.0 * r
    x0 = np.array([[0.5 * r, 0.5 * r, 2.5 * r]], dtype=np.float64)
    v0 = np.array([[0.0, 0.0, -0.5]], dtype=np.float64)
    cfg = DemConfig(
        n=1,
        model_mask=MODEL_CONTACT | MODEL_FACET,
        contact=contact_params(mu=0.5),
        surf_contact=contact_params(mu=0.5),
        facet_vertices=np.array([[-a, -a, 0.0], [-a, a, 0.0], [a, -a, 0.0]]),
        facet_v=(0.0, 0.0, 0.0),
        surf_A=0.0,
        surf_h0=1.0e-9,
    )
    return Scenario("surface_drop", cfg, x0, v0, _zeros(1), _zeros(1))


_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")  # package data: the aggregate of resources/aggregate.h


def aggregate_coordinates() -> np.ndarray:
    """The 150-particle soot aggregate of resources/aggregate.h:8-158, in units of r_part (fixture
    extracted by tests/golden/make_golden.py into libgran_b200/data/aggregate_150.npy)."""
    return np.load(os.path.join(_DATA, "aggregate_150.npy"))


def aggregate(models: int = MODEL_SINTER | MODEL_HAMAKER, neighbor_list: bool = True) -> Scenario:
    """BASELINE config 1: aggregate.h scaled by r_part (Writerside/topics/Custom-step-handler.md:260-264)
    with a deterministic kick: every 10th particle gets v = (0, 0, 5) (style of test/sintered.cpp:88-96)."""
    x0 = aggregate_coordinates() * R_PART
    n = len(x0)
    v0 = _zeros(n)
    v0[::10] = (0.0, 0.0, 5.0)
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=neighbor_list, rng_seed=0)
    return Scenario("aggregate_150", cfg, x0, v0, _zeros(n), _zeros(n))


# ---------------------------------------------------------------------------------------------
# synthetic generators for the larger BASELINE configs
# ---------------------------------------------------------------------------------------------
def _native():
    """The C-ABI library's host-side generators (libgran_b200/csrc/generators.cu); they need no GPU."""
    from .system import load_library

    return load_library()


def fractal_aggregate_coordinates(n: int, seed: int = 12345, r: float = R_PART) -> np.ndarray:
    """Deterministic particle-cluster sticking (configs 2 and 3): each new particle is placed at exactly
    2r from a uniformly chosen existing particle in a uniform random direction and rejected if it comes
    within 2r(1 + 1e-9) of any other particle (lg_gen_fractal_aggregate)."""
    import ctypes as C

    pts = np.zeros((n, 3), dtype=np.float64)
    rc = _native().lg_gen_fractal_aggregate(int(n), int(seed), float(r), pts.ctypes.data_as(C.POINTER(C.c_double)))
    if rc != 0:
        raise RuntimeError(_native().lg_last_error().decode(errors="replace"))
    return pts


def fractal_aggregate(n: int = 10_000, models: int = MODEL_CONTACT | MODEL_HAMAKER, seed: int = 12345,
                      v_seed: int = 1) -> Scenario:
    """BASELINE config 2: N-particle synthetic fractal aggregate, v0 ~ U(-0.5, 0.5) per component."""
    x0 = fractal_aggregate_coordinates(n, seed)
    rng = np.random.Generator(np.random.MT19937(v_seed))
    v0 = rng.uniform(-0.5, 0.5, size=(n, 3))
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=True)
    return Scenario(f"fractal_{n}", cfg, x0, v0, _zeros(n), _zeros(n))


def sintered_aggregate_on_facet(n: int = 100_000, seed: int = 12345) -> Scenario:
    """BASELINE config 3: sintered fractal aggregate resting 1.5 r above a facet that moves +z at 0.5 m/s
    (driver calls facet_update_positions(dt) every step, test/surface.cpp force model with A = 1e-20)."""
    x0 = fractal_aggregate_coordinates(n, seed)
    z_min = x0[:, 2].min()
    lo, hi = x0.min(axis=0), x0.max(axis=0)
    ext = 4.0 * max(hi[0] - lo[0], hi[1] - lo[1]) + 100.0 * R_PART
    cx, cy = 0.5 * (lo[0] + hi[0]), 0.5 * (lo[1] + hi[1])
    zf = z_min - 1.5 * R_PART
    verts = np.array([[cx - ext, cy - ext, zf], [cx - ext, cy + 3.0 * ext, zf], [cx + 3.0 * ext, cy - ext, zf]])
    cfg = DemConfig(
        n=n,
        model_mask=MODEL_SINTER | MODEL_HAMAKER | MODEL_FACET,
        use_neighbor_list=True,
        facet_vertices=verts,
        facet_v=(0.0, 0.0, 0.5),
        surf_A=1.0e-20,
        surf_contact=contact_params(mu=0.5),
    )
    return Scenario(f"sintered_on_facet_{n}", cfg, x0, _zeros(n), _zeros(n), _zeros(n))


def _random_rotations(rng, count: int) -> np.ndarray:
    q = rng.normal(size=(count, 4))
    q /= np.linalg.norm(q, axis=1)[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    rot = np.empty((count, 3, 3))
    rot[:, 0, 0] = 1 - 2 * (y * y + z * z)
    rot[:, 0, 1] = 2 * (x * y - z * w)
    rot[:, 0, 2] = 2 * (x * z + y * w)
    rot[:, 1, 0] = 2 * (x * y + z * w)
    rot[:, 1, 1] = 1 - 2 * (x * x + z * z)
    rot[:, 1, 2] = 2 * (y * z - x * w)
    rot[:, 2, 0] = 2 * (x * z - y * w)
    rot[:, 2, 1] = 2 * (y * z + x * w)
    rot[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return rot


def multi_aggregate_collision(n_copies: int = 6667, models: int = MODEL_SINTER | MODEL_HAMAKER, seed: int = 2,
                              speed: float = 0.5) -> Scenario:
    """BASELINE config 4: translated + randomly rotated copies of aggregate.h on a 3-D grid
    (pitch = bounding-sphere diameter + 6 r), left half moving +x, right half -x (as main.cpp:74,89).
    6667 copies -> 1 000 050 particles.  The grid is elongated in x (2:1:1) so it slab-decomposes."""
    base = aggregate_coordinates() * R_PART
    base = base - base.mean(axis=0)
    radius = np.linalg.norm(base, axis=1).max() + R_PART
    pitch = 2.0 * radius + 6.0 * R_PART
    ny = nz = max(1, int(round((n_copies / 2.0) ** (1.0 / 3.0))))
    nx = int(math.ceil(n_copies / (ny * nz)))
    rng = np.random.Generator(np.random.MT19937(seed))
    rot = _random_rotations(rng, n_copies)
    idx = np.arange(n_copies)
    ix, iy, iz = idx // (ny * nz), (idx // nz) % ny, idx % nz
    centers = np.stack([ix, iy, iz], axis=1).astype(np.float64) * pitch
    x0 = (np.einsum("cij,pj->cpi", rot, base) + centers[:, None, :]).reshape(-1, 3)
    n = len(x0)
    v0 = _zeros(n)
    mid = 0.5 * (x0[:, 0].min() + x0[:, 0].max())
    agg_left = centers[:, 0] < mid
    v0[:, 0] = np.repeat(np.where(agg_left, speed, -speed), len(base))
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=True, rng_seed=0)
    return Scenario(f"multi_aggregate_{n}", cfg, np.ascontiguousarray(x0), v0, _zeros(n), _zeros(n))


def multi_aggregate_in_contact(n_copies: int = 6667, overlap: float = None, models: int = MODEL_SINTER | MODEL_HAMAKER,
                               seed: int = 2, speed: float = 0.5) -> Scenario:
    """Config 4 at the moment of impact: the same rotated copies of aggregate.h on the same grid, but every aggregate is
    slid along -x until it first touches its -x neighbour with an overlap of `overlap` * r (exact first-contact
    distance of the two rigid particle sets), so that whole rows of aggregates are in Luding contact and aggregates of
    neighbouring grid rows come within the Verlet radius.  Default overlap: the static equilibrium k * overlap =
    F_Hamaker(h0) (7e-8 r), where a touching pair neither rebounds nor sinks, so the contacts persist through a run.
    Velocities as config 4 (left half +x, right half -x)."""
    base_scn = multi_aggregate_collision(n_copies, models=models, seed=seed, speed=speed)
    if overlap is None:
        c, r, h = base_scn.cfg, R_PART, base_scn.cfg.hamaker_h0
        f_h = c.hamaker_A / 6.0 * 64.0 * r ** 6 / (h * h * (2.0 * r + h) ** 3 * (4.0 * r + h) ** 2)  # hamaker_force.h:36-40 at h = h0
        overlap = f_h / c.contact[0] / r
    x = base_scn.x0.reshape(n_copies, -1, 3).copy()
    m = x.shape[1]
    ny = nz = max(1, int(round((n_copies / 2.0) ** (1.0 / 3.0))))
    d_touch = (2.0 - overlap) * R_PART
    shift = np.zeros(n_copies)
    # aggregate c sits at grid (ix, iy, iz) with c = (ix * ny + iy) * nz + iz: its -x neighbour is c - ny * nz
    stride = ny * nz
    for c in range(stride, n_copies):
        a = x[c - stride] - np.array([shift[c - stride], 0.0, 0.0])   # neighbour, already slid
        b = x[c]
        dy = b[:, None, 1] - a[None, :, 1]
        dz = b[:, None, 2] - a[None, :, 2]
        lat2 = dy * dy + dz * dz
        hit = lat2 < d_touch * d_touch
        if not hit.any():
            shift[c] = shift[c - stride]
            continue
        dx = b[:, None, 0] - a[None, :, 0]
        slide = np.where(hit, dx - np.sqrt(np.where(hit, d_touch * d_touch - lat2, 0.0)), np.inf)
        shift[c] = slide.min()
    x[:, :, 0] -= shift[:, None]
    x0 = np.ascontiguousarray(x.reshape(-1, 3))

    def make_bonds(unit_vectors):
        # bridges INSIDE the aggregates only: detected on the separated layout (sinter_bridge.h:148-211 via
        # lg_bond_table_create), then carried along with each aggregate's slide; the touching pairs BETWEEN aggregates
        # are closer than critical_separation and would otherwise be sintered together instead of being in contact
        from .system import host_bond_table

        ij, conn, ln = host_bond_table(base_scn.x0, base_scn.cfg.r_part, base_scn.cfg.critical_separation, unit_vectors)
        conn = conn.reshape(-1, 5, 3).copy()
        conn[:, :, 0] -= shift[ij[:, 0] // m][:, None]
        return ij, conn.reshape(-1, 15), ln

    return Scenario(f"multi_aggregate_in_contact_{len(x0)}", base_scn.cfg, x0, base_scn.v0, base_scn.theta0, base_scn.omega0,
                    make_bonds=make_bonds)


def random_packing(n: int = 1_250_000, volume_fraction: float = 0.30, models: int = MODEL_CONTACT | MODEL_HAMAKER,
                   seed: int = 3, aspect: float = 8.0, v_scale: float = 0.05, min_gap_factor: float = 1.0) -> Scenario:
    """BASELINE config 5: random sequential addition of non-overlapping spheres to the given volume
    fraction in a box elongated in x (aspect:1:1), small random velocities (lg_gen_random_packing)."""
    import ctypes as C

    pts = np.empty((n, 3), dtype=np.float64)
    box = np.zeros(3)
    dp = C.POINTER(C.c_double)
    rc = _native().lg_gen_random_packing(int(n), float(volume_fraction), float(aspect), float(min_gap_factor), int(seed),
                                         float(R_PART), pts.ctypes.data_as(dp), box.ctypes.data_as(dp))
    if rc != 0:
        raise RuntimeError(_native().lg_last_error().decode(errors="replace"))
    rng = np.random.Generator(np.random.MT19937(seed))
    v0 = rng.uniform(-v_scale, v_scale, size=(n, 3))
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=True)
    return Scenario(f"random_packing_{n}", cfg, pts, v0, _zeros(n), _zeros(n))


def lattice_block(nx: int, ny: int, nz: int, models: int = MODEL_CONTACT | MODEL_HAMAKER, jitter: float = 0.02,
                  spacing: float = 2.0, seed: int = 7, v_scale: float = 0.5) -> Scenario:
    """Cheap dense test input: a jittered simple-cubic block (the 10x10x10 / 22^3 probes of BASELINE.md section 2)."""
    g = np.stack(np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij"), axis=-1).reshape(-1, 3)
    rng = np.random.Generator(np.random.MT19937(seed))
    x0 = (g.astype(np.float64) * spacing + rng.uniform(-jitter, jitter, size=g.shape)) * R_PART
    n = len(x0)
    v0 = rng.uniform(-v_scale, v_scale, size=(n, 3))
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=True)
    return Scenario(f"lattice_{nx}x{ny}x{nz}", cfg, np.ascontiguousarray(x0), v0, _zeros(n), _zeros(n))


def stress_block(nx: int = 6, ny: int = 6, nz: int = 5, models: int = MODEL_CONTACT | MODEL_HAMAKER | MODEL_FACET,
                 neighbor_list: bool = True, seed: int = 11, spacing: float = 1.96, jitter: float = 0.05,
                 v_scale: float = 5.0, w_scale: float = 3.0e8, facet_v=(2.0, -1.0, 0.3)) -> Scenario:
    """Branch-coverage input (not a reference set-up): an over-dense jittered block with large random
    linear and angular velocities so that within a few hundred steps pairs are created, slide (dynamic
    friction), stick, roll, twist and separate; the bottom layer overlaps a tangentially moving facet whose
    edge runs under the block so that the plane, edge and vertex branches of the closest-point search are hit."""
    g = np.stack(np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij"), axis=-1).reshape(-1, 3)
    rng = np.random.Generator(np.random.MT19937(seed))
    x0 = (g.astype(np.float64) * spacing + rng.uniform(-jitter, jitter, size=g.shape)) * R_PART
    n = len(x0)
    v0 = rng.uniform(-v_scale, v_scale, size=(n, 3))
    w0 = rng.uniform(-w_scale, w_scale, size=(n, 3))
    lx, ly = (nx - 1) * spacing * R_PART, (ny - 1) * spacing * R_PART
    zf = -0.9 * R_PART
    # right triangle covering roughly half of the block footprint: hypotenuse crosses under the block
    verts = np.array([[-2.0 * R_PART, -2.0 * R_PART, zf], [-2.0 * R_PART, ly + 2.0 * R_PART, zf],
                      [lx + 2.0 * R_PART, -2.0 * R_PART, zf]])
    cfg = DemConfig(n=n, model_mask=models, use_neighbor_list=neighbor_list, facet_vertices=verts,
                    facet_v=tuple(facet_v), surf_A=1.0e-20, surf_contact=contact_params(mu=0.5), rebuild_every=5)
    return Scenario(f"stress_{nx}x{ny}x{nz}", cfg, np.ascontiguousarray(x0), v0, _zeros(n), w0)


def aggregates_with_dense_block(n_copies: int = 20, block: int = 7, spacing: float = 1.96, jitter: float = 0.04,
                                models: int = MODEL_SINTER | MODEL_HAMAKER, seed: int = 5, v_scale: float = 2.0) -> Scenario:
    """Branch-coverage input for the short-row step kernel (not a reference set-up): `n_copies` soot aggregates (rows of
    ~10 neighbors) followed by one over-dense jittered block of block^3 particles (rows of ~80 neighbors, overlapping
    non-bonded pairs, six or more bonds per particle when the sinter model is on).  The mean row stays short, so the
    production dispatch picks the row kernel while single rows exceed 64 pairs and single CTAs hold several hundred
    bond records."""
    agg = multi_aggregate_collision(n_copies, models=models)
    rng = np.random.Generator(np.random.MT19937(seed))
    g = np.stack(np.meshgrid(np.arange(block), np.arange(block), np.arange(block), indexing="ij"), axis=-1).reshape(-1, 3)
    xb = (g.astype(np.float64) * spacing + rng.uniform(-jitter, jitter, size=g.shape)) * R_PART
    xb += np.array([agg.x0[:, 0].max() + 12.0 * R_PART, 0.0, 0.0])
    x0 = np.ascontiguousarray(np.concatenate([agg.x0, xb]))
    n = len(x0)
    v0 = np.concatenate([agg.v0, rng.uniform(-v_scale, v_scale, size=xb.shape)])
    cfg = agg.cfg.with_(n=n)
    return Scenario(f"aggregates_{n_copies}_block_{block}", cfg, x0, np.ascontiguousarray(v0), _zeros(n), _zeros(n))
