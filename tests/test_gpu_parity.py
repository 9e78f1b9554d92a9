"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the committed reference outputs.

Bars (BASELINE.json north_star): neighbor / contact sets bit-exact; per-step forces and torques within
1e-10 relative -- the CUDA path is in fact BIT-IDENTICAL in x, v, theta, omega, a, alpha at every compared
step because it keeps the reference's operand and accumulation order (compiled with --fmad=false); energies
over 1000 steps therefore match exactly as well (tolerance written in test_energy_1000_steps: 1e-12 relative
for the device tree reductions vs the reference's serial sums)."""
import numpy as np
import pytest

from libgran_b200 import scenarios as S

from helpers import FIELDS, advance, assert_fields_equal, golden, golden_scenarios, unit_vectors_for

pytestmark = pytest.mark.gpu


def make_gpu(scn, oracle, variant=None, max_contacts=None):
    from libgran_b200.system import from_config

    g = from_config(scn.cfg, *scn.state, unit_vectors=unit_vectors_for(scn.cfg, oracle), device=0, variant=variant)
    return g


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("name", sorted(golden_scenarios()))
def test_gpu_matches_reference_golden(lib, oracle, name, variant):
    """CUDA path vs outputs of the unmodified reference (tests/golden), bit for bit."""
    scn, moves = golden_scenarios()[name]
    g = golden(name)
    gpu = make_gpu(scn, oracle, variant)
    done = 0
    for cp in g["checkpoints"]:
        advance(gpu, scn.cfg, done, int(cp), moves)
        done = int(cp)
        for f in FIELDS:
            got, want = gpu.get(f), g[f"{f}_{cp}"]
            assert np.array_equal(got, want), (
                f"{name} v{variant}: {f} at step {cp}: max rel "
                f"{np.abs(got - want).max() / max(np.abs(want).max(), 1e-300):.3e}")
    rp, rc = gpu.rows()
    assert np.array_equal(rp, g["row_ptr"]) and np.array_equal(rc, g["cols"]), f"{name}: neighbor rows"


LIVE = [
    ("slab h nl", lambda: S.slab_and_projectile(S.MODEL_HAMAKER, True), 100, False),
    ("lattice s+h nl", lambda: S.sintered_lattice(S.MODEL_SINTER | S.MODEL_HAMAKER, True), 300, False),
    ("aggregate s allpairs", lambda: S.aggregate(S.MODEL_SINTER, False), 60, False),
    ("stress s+f allpairs", lambda: S.stress_block(models=S.MODEL_SINTER | S.MODEL_FACET, neighbor_list=False), 100, True),
    ("stress c allpairs", lambda: S.stress_block(models=S.MODEL_CONTACT, neighbor_list=False), 100, False),
    ("stress c+h+f nl big", lambda: S.stress_block(14, 14, 10, models=S.MODEL_CONTACT | S.MODEL_HAMAKER | S.MODEL_FACET), 120, True),
    ("stress s+h+f nl big", lambda: S.stress_block(14, 14, 10, models=S.MODEL_SINTER | S.MODEL_HAMAKER | S.MODEL_FACET), 120, True),
    ("lattice 16^3 c+h nl", lambda: S.lattice_block(16, 16, 16), 60, False),
    ("fractal 3000 s+h nl", lambda: S.fractal_aggregate(3000, models=S.MODEL_SINTER | S.MODEL_HAMAKER), 100, False),
    ("multi aggregate 20 s+h nl", lambda: S.multi_aggregate_collision(20), 100, False),
    ("packing 4000 c+h nl", lambda: S.random_packing(4000, volume_fraction=0.3, v_scale=3.0), 100, False),
]


@pytest.mark.parametrize("variant", [0, 1])
@pytest.mark.parametrize("case", LIVE, ids=[c[0] for c in LIVE])
def test_gpu_matches_oracle_lockstep(lib, oracle, case, variant):
    _, make, steps, moves = case
    scn = make()
    gpu = make_gpu(scn, oracle, variant)
    cpu = oracle.OracleSystem(scn.cfg, *scn.state)
    chunk = max(1, steps // 5)
    for s in range(0, steps, chunk):
        stop = min(steps, s + chunk)
        advance(gpu, scn.cfg, s, stop, moves)
        advance(cpu, scn.cfg, s, stop, moves)
        assert_fields_equal(gpu, cpu, f"{scn.name} v{variant} step {stop}")
        if scn.cfg.use_neighbor_list:
            (gp, gc), (cp, cc) = gpu.rows(), cpu.rows()
            assert np.array_equal(gp, cp) and np.array_equal(gc, cc), f"{scn.name}: neighbor rows at step {stop}"
    c = gpu.counters()
    assert c["contacts"] == cpu.n_contacts(), "contact set size"
    assert c["bonds"] == cpu.n_bonds(), "bond set size"
    # hidden per-pair state
    rp, rc = cpu.rows()
    rng = np.random.default_rng(1)
    n = scn.cfg.n
    for i in rng.integers(0, n, size=40):
        i = int(i)
        for j in rc[rp[i]:rp[i + 1]][:20]:
            j = int(j)
            assert np.array_equal(gpu.contact_springs(i, j), cpu.contact_springs(i, j)), f"springs ({i},{j})"
            gb, cb = gpu.bond(i, j), cpu.bond(i, j)
            assert (gb is None) == (cb is None)
            if gb is not None:
                assert np.array_equal(gb[0], cb[0]) and np.array_equal(gb[1], cb[1]), f"connectors ({i},{j})"


def test_neighbor_rows_ragged_and_empty(lib, oracle):
    """Edge cases of the rebuild: isolated particles (empty rows), a single particle, everything in one cell,
    a pair sitting numerically at the Verlet radius."""
    r = S.R_PART
    rv = 5.0 * r
    x = np.array([[0, 0, 0], [rv, 0, 0], [np.nextafter(rv, 0), 0, 0], [100 * r, 0, 0], [0.3 * r, 0.1 * r, 0], [0, 3 * r, 4 * r],
                  [0, 3 * r, np.nextafter(4 * r, 0)]], dtype=np.float64)
    n = len(x)
    z = np.zeros((n, 3))
    cfg = S.DemConfig(n=n, model_mask=S.MODEL_HAMAKER, use_neighbor_list=True)
    from libgran_b200.system import from_config

    gpu = from_config(cfg, x, z, z, z, device=0)
    cpu = oracle.OracleSystem(cfg, x, z, z, z)
    assert gpu.n_pairs() == 0  # empty until the first update (rotational_binary_system_neighbors_omp.h:66)
    gpu.update_neighbor_list()
    cpu.update_neighbor_list()
    (gp, gc), (cp, cc) = gpu.rows(), cpu.rows()
    assert np.array_equal(gp, cp) and np.array_equal(gc, cc)
    assert gp[4] - gp[3] == 0  # the isolated particle has an empty row
    one = from_config(cfg.with_(n=1), x[:1], z[:1], z[:1], z[:1], device=0)
    one.update_neighbor_list()
    one.do_step(cfg.dt)
    assert one.n_pairs() == 0 and np.all(one.get_a() == 0.0)


def test_neighbor_rows_large_vs_oracle(lib, oracle):
    """100 k particles: pair set identical to the oracle's (grid search, itself pinned to brute force)."""
    scn = S.lattice_block(50, 50, 40, jitter=0.3, spacing=2.3)
    from libgran_b200.system import from_config

    gpu = from_config(scn.cfg, *scn.state, device=0)
    cpu = oracle.OracleSystem(scn.cfg, *scn.state)
    gpu.update_neighbor_list()
    cpu.update_neighbor_list()
    (gp, gc), (cp, cc) = gpu.rows(), cpu.rows()
    assert np.array_equal(gp, cp) and np.array_equal(gc, cc)
    # size-independent properties: symmetric, sorted, no self pairs
    rows = [gc[gp[i]:gp[i + 1]] for i in range(0, scn.cfg.n, 997)]
    for k, row in enumerate(rows):
        i = k * 997
        assert np.all(np.diff(row) > 0) and i not in row
        for j in row[:5]:
            assert i in gc[gp[j]:gp[j + 1]]


def test_energy_1000_steps(lib, oracle):
    """compute_energy over 1000 steps (north_star): device tree reductions vs the serial sums of
    test/compute_energy.cpp on a bit-identical trajectory; tolerance 1e-12 relative."""
    scn = S.aggregate()
    gpu = make_gpu(scn, oracle)
    cpu = oracle.OracleSystem(scn.cfg, *scn.state)
    for stop in range(100, 1001, 100):
        advance(gpu, scn.cfg, stop - 100, stop, False)
        advance(cpu, scn.cfg, stop - 100, stop, False)
        d = gpu.diagnostics()
        o = oracle.oracle_diagnostics(cpu.get("x"), cpu.get("v"), cpu.get("omega"), scn.cfg.mass, scn.cfg.inertia)
        for key in ("ke", "p", "rg"):
            assert abs(d[key] - o[key]) <= 1e-12 * abs(o[key]), (stop, key, d[key], o[key])
        assert np.allclose(d["com"], o["com"], rtol=1e-12, atol=1e-22)
        assert abs(d["l"] - o["l"]) <= 1e-9 * max(abs(o["l"]), 1e-40)
    assert_fields_equal(gpu, cpu, "aggregate after 1000 steps")


def test_deterministic_run_to_run(lib, oracle):
    scn = S.stress_block(10, 10, 8, models=S.MODEL_SINTER | S.MODEL_HAMAKER | S.MODEL_FACET)
    outs = []
    for _ in range(2):
        gpu = make_gpu(scn, oracle)
        advance(gpu, scn.cfg, 0, 80, True)
        outs.append({f: gpu.get(f) for f in FIELDS})
    for f in FIELDS:
        assert np.array_equal(outs[0][f], outs[1][f])


def test_contact_capacity_error_is_loud(lib, oracle):
    from libgran_b200.system import LgError, from_config

    scn = S.stress_block(models=S.MODEL_CONTACT)
    gpu = from_config(scn.cfg, *scn.state, device=0)
    gpu2 = None
    with pytest.raises(LgError) as ei:
        from libgran_b200.system import GranularSystem

        gpu2 = GranularSystem(*scn.state, scn.cfg.r_part, scn.cfg.mass, scn.cfg.inertia, device=0, r_verlet=scn.cfg.r_verlet)
        gpu2.set_max_contacts(1)
        gpu2.set_contact(scn.cfg.contact)
        gpu2.run(scn.cfg.dt, 20, 5)
        gpu2.synchronize()
    assert ei.value.code == -6


def test_set_bonds_explicit_tables(lib, oracle):
    """lg_set_bonds with the tables read back from the oracle == lg_build_sinter_bonds."""
    scn = S.sintered_lattice(S.MODEL_SINTER | S.MODEL_HAMAKER, True)
    cfg = scn.cfg
    cpu = oracle.OracleSystem(cfg, *scn.state)
    ij, conn, rl = [], [], []
    for i in range(cfg.n):
        for j in range(cfg.n):
            b = cpu.bond(i, j)
            if b is not None:
                ij.append((i, j)); conn.append(b[0]); rl.append(b[1])
    from libgran_b200.system import GranularSystem

    gpu = GranularSystem(*scn.state, cfg.r_part, cfg.mass, cfg.inertia, device=0, r_verlet=cfg.r_verlet)
    gpu.set_bonds(cfg.contact, cfg.k_d, cfg.gamma_d, ij, conn, rl)
    gpu.set_hamaker(cfg.hamaker_A, cfg.hamaker_h0)
    advance(gpu, cfg, 0, 100, False)
    advance(cpu, cfg, 0, 100, False)
    assert_fields_equal(gpu, cpu, "explicit bond tables")


def test_do_step_and_run_are_equivalent(lib, oracle):
    scn = S.slab_and_projectile()
    a = make_gpu(scn, oracle)
    b = make_gpu(scn, oracle)
    a.run(scn.cfg.dt, 50, 20)
    for s in range(50):
        if s % 20 == 0:
            b.update_neighbor_list()
        b.do_step(scn.cfg.dt)
    assert_fields_equal(a, b, "run vs do_step loop")
