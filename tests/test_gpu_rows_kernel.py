"""The kernel bench.py times -- step_rows_kernel, LG_MATH_FAST -- checked DIRECTLY against the CPU oracle
(oracle.OracleSystem, itself bit-identical to the unmodified reference, tests/test_oracle.py).

Tolerance (BASELINE.json north_star): per-step forces and torques within 1e-10 relative, on identical inputs
("teacher forcing": both sides are constructed from the same state bits).  `rel_err` is relative PER PARTICLE VECTOR,
|a_gpu - a_ref| / max(|a_ref|, floor) with floor = 1e-6 of the largest |a_ref| of the system, so that particles whose
forces cancel to (almost) zero are measured against the system's force scale instead of against rounding noise.
Torques of sinter bridges need one more, conditioning-aware term.  A bridge's four off-axis springs (sinter_bridge.h:185-206)
pull with equal forces at opposite arms +-ds r, +-db r, so their torques (c_i - x_i) x dF cancel in exact arithmetic and
alpha_i is the RESIDUAL of terms up to 1e9 times larger (aggregate.h's touching spheres sit 2.9e-6 r inside the rest length:
each spring carries 4e-10 N, each torque term 3.7e18 rad/s^2, the sum ~4e9).  The reference's own result therefore carries
a rounding error of eps * sum|terms|; two correct implementations cannot agree better than that, so the comparison allows
    |alpha_gpu - alpha_ref| <= 1e-10 * max(|alpha_ref|, floor) + 4 eps * sum_bonds 4 r k_d |dist - 2 r| / I
(`torque_term_scale`, computed from the positions).  In LG_MATH_EXACT the kernels reproduce the reference's bits, terms
and all (tests/test_gpu_parity.py); this allowance only applies to the relaxed-arithmetic mode.

Both instantiations of the kernel are covered on every case:
  * force-only (MODE_FORCE): do_step(0) leaves the closing evaluation pending, get_a() flushes it through the row
    kernel at exactly the constructor's x, v, omega;
  * fused (MODE_FUSED): do_step(0), do_step(0), do_step(dt) -- the third call is ONE fused launch (evaluation at the
    unchanged state + integration), so a = (v2 - v0) / dt and alpha = (omega2 - omega0) / dt; the comparison is made
    on the increments, with 2 ulp of v added for the subtraction.
Cases are chosen to reach every code path of the kernel: rows longer than 64 pairs (the marked-pair mask ends),
more than RBC = 288 bond records in one CTA (bridges evaluated from global memory), more than RPC = 1536 pairs in
one CTA (columns read from global memory), overlapping non-bonded pairs (Luding contact out of line), facets."""
import numpy as np
import pytest

from libgran_b200 import scenarios as S

from helpers import unit_vectors_for

pytestmark = pytest.mark.gpu

FORCE_RTOL = 1e-10

# name -> (scenario factory, kernel variant: 1 = production dispatch (must pick the row kernel), 2 = row kernel forced)
CASES = {
    "aggregate s+h": (lambda: S.aggregate(), 1),
    "cfg4 slice 100 aggregates s+h": (lambda: S.multi_aggregate_collision(100), 1),
    "aggregates + dense block s+h": (lambda: S.aggregates_with_dense_block(20), 1),
    "aggregates + dense block c+h": (lambda: S.aggregates_with_dense_block(20, models=S.MODEL_CONTACT | S.MODEL_HAMAKER), 1),
    "fractal 3000 s+h forced": (lambda: S.fractal_aggregate(3000, models=S.MODEL_SINTER | S.MODEL_HAMAKER), 2),
    "fractal 3000 c+h forced": (lambda: S.fractal_aggregate(3000), 2),
    "stress block s+h forced": (lambda: S.stress_block(14, 14, 10, models=S.MODEL_SINTER | S.MODEL_HAMAKER), 2),
    "stress block s+h+f forced": (lambda: S.stress_block(14, 14, 10, models=S.MODEL_SINTER | S.MODEL_HAMAKER | S.MODEL_FACET), 2),
    "stress block c forced": (lambda: S.stress_block(10, 10, 8, models=S.MODEL_CONTACT), 2),
    "packing 4000 c+h forced": (lambda: S.random_packing(4000, volume_fraction=0.3, v_scale=300.0), 2),
}


def rel_err(a, b, slack=0.0):
    """max over particles of (|a - b| - slack) / max(|b|, 1e-6 max|b|); slack: per-particle absolute allowance."""
    nb = np.linalg.norm(b, axis=1)
    floor = max(nb.max(), 1e-300) * 1e-6
    return float((np.maximum(np.linalg.norm(a - b, axis=1) - slack, 0.0) / np.maximum(nb, floor)).max())


def torque_term_scale(cfg, x):
    """4 eps * sum over the bridges of particle i of the four cancelling torque terms r * k_d * |dist - 2 r| / I: the rounding
    error the reference's own alpha carries (module docstring).  Bridges are what the sinter constructor builds from x:
    every pair with |dist - 2 r| <= critical_separation (sinter_bridge.h:155-157).  Zero without the sinter model."""
    n = len(x)
    out = np.zeros(n)
    if not (cfg.model_mask & S.MODEL_SINTER):
        return out
    from scipy.spatial import cKDTree

    pairs = cKDTree(x).query_pairs(2.0 * cfg.r_part + cfg.critical_separation * (1 + 1e-9), output_type="ndarray")
    if len(pairs) == 0:
        return out
    d = np.linalg.norm(x[pairs[:, 0]] - x[pairs[:, 1]], axis=1)
    st = np.abs(d - 2.0 * cfg.r_part)
    keep = st <= cfg.critical_separation
    term = 4.0 * cfg.r_part * cfg.k_d * st[keep] / cfg.inertia
    np.add.at(out, pairs[keep, 0], term)
    np.add.at(out, pairs[keep, 1], term)
    return 4.0 * np.finfo(np.float64).eps * out


def make_gpu(scn, oracle, state, variant):
    from libgran_b200.system import from_config

    g = from_config(scn.cfg, *state, unit_vectors=unit_vectors_for(scn.cfg, oracle), device=0, variant=variant, math_mode=1)
    return g


def check_state(scn, oracle, state, variant, what):
    cfg = scn.cfg
    gpu = make_gpu(scn, oracle, state, variant)
    cpu = oracle.OracleSystem(cfg, *state)
    if cfg.use_neighbor_list:
        gpu.update_neighbor_list()
        cpu.update_neighbor_list()
        (gp, gc), (cp, cc) = gpu.rows(), cpu.rows()
        assert np.array_equal(gp, cp) and np.array_equal(gc, cc), f"{what}: neighbor rows"
    # ---- force-only instantiation ---------------------------------------------------------------------------------
    gpu.do_step(0.0)
    cpu.do_step(0.0)
    ga, gal = gpu.get("a"), gpu.get("alpha")
    ca, cal = cpu.get("a"), cpu.get("alpha")
    tscale = torque_term_scale(cfg, np.asarray(state[0]))
    assert rel_err(ga, ca) <= FORCE_RTOL, f"{what}: a (force-only row kernel) rel err {rel_err(ga, ca):.3e}"
    if np.abs(cal).max() > 0:
        assert rel_err(gal, cal, tscale) <= FORCE_RTOL, f"{what}: alpha (force-only row kernel) rel err {rel_err(gal, cal, tscale):.3e}"
    assert gpu.last_step_kernel().startswith("step_rows"), f"{what}: dispatch picked {gpu.last_step_kernel()}"
    assert gpu.counters()["contacts"] == cpu.n_contacts(), f"{what}: contact set size"
    # ---- fused instantiation -----------------------------------------------------------------------------------------
    for s in (gpu, cpu):
        s.do_step(0.0)      # a, alpha are current -> plain integration with dt = 0, closing evaluation left pending
    v0, w0 = gpu.get("v"), gpu.get("omega")
    assert np.array_equal(v0, cpu.get("v")) and np.array_equal(w0, cpu.get("omega"))
    dt = cfg.dt
    for s in (gpu, cpu):
        s.do_step(dt)       # GPU: one fused launch = evaluation at the unchanged state + integration
    dv_g, dw_g = gpu.get("v") - v0, gpu.get("omega") - w0
    dv_c, dw_c = cpu.get("v") - v0, cpu.get("omega") - w0
    for name, dg, dc, base, cond in (("v", dv_g, dv_c, v0, 0.0), ("omega", dw_g, dw_c, w0, tscale * dt)):
        nb = np.linalg.norm(dc, axis=1)
        floor = max(nb.max(), 1e-300) * 1e-6
        slack = 2.0 * np.spacing(np.abs(base).max(axis=1) + nb) + cond   # the subtraction v2 - v0 itself (+ torque conditioning)
        err = np.linalg.norm(dg - dc, axis=1)
        bad = err > FORCE_RTOL * np.maximum(nb, floor) + slack
        assert not bad.any(), f"{what}: increment of {name} (fused row kernel): worst {float((err / np.maximum(nb, floor)).max()):.3e}"
    assert gpu.last_step_kernel().startswith("step_rows")
    gpu.close()
    cpu.close()


@pytest.mark.parametrize("name", sorted(CASES))
def test_row_kernel_forces_vs_oracle(lib, oracle, name):
    make, variant = CASES[name]
    scn = make()
    cfg = scn.cfg
    # teacher-forcing states: points of the ORACLE's trajectory
    drv = oracle.OracleSystem(cfg, *scn.state)
    done = 0
    for cp in (0, 40, 160):
        drv.run(cfg.dt, cp - done, cfg.rebuild_every if cfg.use_neighbor_list else 0)
        done = cp
        state = tuple(drv.get(f) for f in ("x", "v", "theta", "omega"))
        check_state(scn, oracle, state, variant, f"{name} @ step {cp}")
    drv.close()


def test_row_kernel_is_what_runs(lib, oracle):
    """The production dispatch (variant 1, fast math) launches step_rows_kernel on short-row inputs: its per-launch
    profile counts one launch per evaluation and the library reports the kernel's name."""
    scn = S.multi_aggregate_collision(20)
    g = make_gpu(scn, oracle, scn.state, 1)
    g.update_neighbor_list()
    g.do_step(scn.cfg.dt)
    assert g.last_step_kernel().startswith("step_rows")
    g.get("a")   # flush: the force-only launch is the row kernel as well
    assert g.last_step_kernel().startswith("step_rows")
    g.close()
    e = from_exact(scn, oracle)
    e.update_neighbor_list()
    e.do_step(scn.cfg.dt)
    assert e.last_step_kernel() == "step_tiled_kernel"
    e.close()


def from_exact(scn, oracle):
    from libgran_b200.system import from_config

    return from_config(scn.cfg, *scn.state, unit_vectors=unit_vectors_for(scn.cfg, oracle), device=0)


def test_cfg4_full_size_vs_oracle(lib, oracle):
    """BASELINE config 4 at FULL size (6667 aggregates, 1 000 050 particles), the benched kernel in the benched mode:
    neighbor rows bit-equal to the oracle's grid search, bond count equal, a and alpha within 1e-10 of the oracle's
    (the sparse-state oracle scales; the reference's dense tables would need 232 TB), and the size-independent
    properties: rows sorted and symmetric, and -- aggregates are farther apart than r_verlet -- every aggregate's
    internal forces sum to zero (Newton's third law; sum m a over 150 particles vs sum |m a|)."""
    scn = S.multi_aggregate_collision(6667)
    cfg = scn.cfg
    n = cfg.n
    assert n == 1_000_050
    gpu = make_gpu(scn, oracle, scn.state, 1)
    cpu = oracle.OracleSystem(cfg, *scn.state)
    gpu.update_neighbor_list()
    cpu.update_neighbor_list()
    (gp, gc), (cp, cc) = gpu.rows(), cpu.rows()
    assert np.array_equal(gp, cp) and np.array_equal(gc, cc)
    assert gpu.counters()["bonds"] == cpu.n_bonds()
    # sorted, no self pairs, symmetric (pair (i, j) present <=> (j, i) present): compare the sorted edge lists
    i_of = np.repeat(np.arange(n, dtype=np.int64), np.diff(gp))
    assert np.all(gc != i_of)
    same_row = i_of[1:] == i_of[:-1]
    assert np.all(np.diff(gc)[same_row] > 0)
    fwd = i_of * n + gc
    rev = np.sort(gc * n + i_of)
    assert np.array_equal(fwd, rev)   # fwd is already sorted (rows ascending, columns ascending)
    gpu.do_step(0.0)
    cpu.do_step(0.0)
    ga, gal, ca, cal = gpu.get("a"), gpu.get("alpha"), cpu.get("a"), cpu.get("alpha")
    tscale = torque_term_scale(cfg, scn.x0)
    assert rel_err(ga, ca) <= FORCE_RTOL, f"a rel err {rel_err(ga, ca):.3e}"
    assert rel_err(gal, cal, tscale) <= FORCE_RTOL, f"alpha rel err {rel_err(gal, cal, tscale):.3e}"
    per_agg = ga.reshape(-1, 150, 3)
    net = np.linalg.norm(per_agg.sum(axis=1), axis=1)
    tot = np.linalg.norm(per_agg, axis=2).sum(axis=1)
    assert (net <= 1e-9 * tot).all(), f"worst net/total force of an aggregate {float((net / tot).max()):.3e}"
    # the fused launch at full size, against the oracle's increments
    for s in (gpu, cpu):
        s.do_step(0.0)
    v0 = gpu.get("v")
    for s in (gpu, cpu):
        s.do_step(cfg.dt)
    dv_g, dv_c = gpu.get("v") - v0, cpu.get("v") - v0
    nb = np.linalg.norm(dv_c, axis=1)
    err = np.linalg.norm(dv_g - dv_c, axis=1)
    floor = nb.max() * 1e-6
    assert (err <= FORCE_RTOL * np.maximum(nb, floor) + 2.0 * np.spacing(np.abs(v0).max(axis=1) + nb)).all()
    gpu.close()
    cpu.close()
