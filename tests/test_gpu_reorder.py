"""Spatial renumbering (lg_set_reorder): invisible at the C ABI.  A renumbered system must answer in the caller's particle
order -- state, neighbor rows (bit-exact sets), bond tables (same bridges, same connectors: the sinter scan still consumes the
unit-vector generator in the caller's lexicographic pair order), history springs -- and its forces may differ from the
caller-order system only by the summation order inside a row (<= 1e-10 relative, the tolerance north_star states)."""
import numpy as np
import pytest

from libgran_b200 import scenarios as S

from helpers import unit_vectors_for

pytestmark = pytest.mark.gpu


def shuffled(scn, seed=3):
    """The same physical system with its particle ids shuffled: no spatial locality left in the caller's order."""
    rng = np.random.default_rng(seed)
    p = rng.permutation(scn.cfg.n)
    return S.Scenario(scn.name + "_shuffled", scn.cfg, *(np.ascontiguousarray(a[p]) for a in scn.state))


def rel_err(got, want):
    scale = np.maximum(np.linalg.norm(want, axis=1), 1e-6 * np.linalg.norm(want, axis=1).max())
    return float((np.linalg.norm(got - want, axis=1) / scale).max())


CASES = [
    ("packing 6000 c+h", lambda: shuffled(S.random_packing(6000, volume_fraction=0.3, v_scale=3.0)), 1),
    ("fractal 5000 s+h", lambda: shuffled(S.fractal_aggregate(5000, models=S.MODEL_SINTER | S.MODEL_HAMAKER)), 1),
    ("multi aggregate 30 s+h", lambda: shuffled(S.multi_aggregate_collision(30)), 1),
    ("lattice 12^3 c+h dense", lambda: shuffled(S.lattice_block(12, 12, 12, spacing=1.98, jitter=0.02, v_scale=0.05)), 1),
    ("packing 6000 c+h exact", lambda: shuffled(S.random_packing(6000, volume_fraction=0.3, v_scale=3.0)), 0),
]


@pytest.mark.parametrize("name,make,mode", CASES, ids=[c[0] for c in CASES])
def test_reordered_system_answers_in_caller_order(lib, oracle, name, make, mode):
    from libgran_b200.system import from_config

    scn = make()
    cfg = scn.cfg
    uv = unit_vectors_for(cfg, oracle)
    plain = from_config(cfg, *scn.state, unit_vectors=uv, device=0, math_mode=mode)
    reo = from_config(cfg, *scn.state, unit_vectors=uv, device=0, math_mode=mode, reorder=True)
    # state round trip and bond tables before any step
    for f in ("x", "v", "theta", "omega"):
        assert np.array_equal(reo.get(f), plain.get(f)), f"{name}: {f} round trip"
    assert reo.counters()["bonds"] == plain.counters()["bonds"]
    for s in (plain, reo):
        s.update_neighbor_list()
    (pp, pc), (rp, rc) = plain.rows(), reo.rows()
    assert np.array_equal(pp, rp) and np.array_equal(pc, rc), f"{name}: neighbor rows"
    rng = np.random.default_rng(0)
    for i in rng.integers(0, cfg.n, size=60):
        i = int(i)
        for j in pc[pp[i]:pp[i + 1]][:12]:
            a, b = plain.bond(i, int(j)), reo.bond(i, int(j))
            assert (a is None) == (b is None), f"{name}: bond flag ({i},{j})"
            if a is not None:
                assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), f"{name}: connectors ({i},{j})"
    # teacher-forced forces
    for s in (plain, reo):
        s.do_step(0.0)
    assert rel_err(reo.get("a"), plain.get("a")) <= 1e-10, f"{name}: a"
    al = plain.get("alpha")
    if np.abs(al).max() > 0:
        assert rel_err(reo.get("alpha"), al) <= 1e-8, f"{name}: alpha"  # torque sums cancel: conditioning, see test_gpu_rows_kernel
    # a short trajectory incl. rebuilds: same rows, same contact set, positions within rounding growth
    for s in (plain, reo):
        s.run(cfg.dt, 60, cfg.rebuild_every)
    (pp, pc), (rp, rc) = plain.rows(), reo.rows()
    assert np.array_equal(pp, rp) and np.array_equal(pc, rc), f"{name}: neighbor rows after 60 steps"
    assert reo.counters()["contacts"] == plain.counters()["contacts"], f"{name}: contact set size"
    dx = np.abs(reo.get("x") - plain.get("x")).max()
    assert dx <= 1e-9 * cfg.r_part, f"{name}: x drift {dx:.3e}"
    # history springs of pairs in contact answer under the caller's ids
    hits = 0
    for i in range(cfg.n):
        for j in pc[pp[i]:pp[i + 1]]:
            sp = plain.contact_springs(i, int(j))
            if np.any(sp != 0.0):
                sr = reo.contact_springs(i, int(j))
                assert np.allclose(sr, sp, rtol=1e-6, atol=1e-30), f"{name}: springs ({i},{j})"
                hits += 1
                break
        if hits >= 10:
            break


def test_reorder_guards(lib):
    from libgran_b200.system import LgError, from_config

    scn = S.random_packing(500, volume_fraction=0.2)
    g = from_config(scn.cfg, *scn.state, device=0)
    with pytest.raises(LgError):
        g._check(g.lib.lg_set_reorder(g.h, 1))  # after the first lg_set_state
    r = from_config(scn.cfg, *scn.state, device=0, reorder=True)
    with pytest.raises(LgError):
        r.device_ptr("x")


# ---- device-side bond detection (lg_bond_table_create_device) -----------------------------------------------------------
BOND_CASES = [
    ("aggregate 150", lambda: S.aggregate()),
    ("multi aggregate 40", lambda: S.multi_aggregate_collision(40)),
    ("fractal 20000", lambda: S.fractal_aggregate(20000, models=S.MODEL_SINTER | S.MODEL_HAMAKER)),
    ("stress block", lambda: S.stress_block(14, 14, 10, models=S.MODEL_SINTER | S.MODEL_HAMAKER)),
]


@pytest.mark.parametrize("name,make", BOND_CASES, ids=[c[0] for c in BOND_CASES])
def test_device_bond_table_equals_host_scan(lib, oracle, name, make):
    """Same bond set, same connectors and rest lengths, bit for bit: the device scan hands unit vector k to the k-th bond in
    lexicographic (i < j) order, exactly as the reference's constructor consumes its generator."""
    from libgran_b200 import system as sysm

    scn = make()
    cfg = scn.cfg
    uv = unit_vectors_for(cfg, oracle, count=4 * cfg.n + 64)
    h_ij, h_conn, h_len = sysm.host_bond_table(scn.x0, cfg.r_part, cfg.critical_separation, uv)
    d_ij, d_conn, d_len = sysm.device_bond_table(scn.x0, cfg.r_part, cfg.critical_separation, uv, device=0)
    assert len(h_ij) > 0
    ho = np.lexsort((h_ij[:, 1], h_ij[:, 0]))  # the host table lists (i, j), (j, i) per bond; the device one is in CSR order
    do = np.lexsort((d_ij[:, 1], d_ij[:, 0]))
    assert np.array_equal(h_ij[ho], d_ij[do]), f"{name}: bond set"
    assert np.array_equal(h_conn[ho], d_conn[do]), f"{name}: connectors"
    assert np.array_equal(h_len[ho], d_len[do]), f"{name}: rest lengths"


def test_device_bond_table_collinear_draw_falls_back(lib, oracle):
    """A unit vector collinear with its bond axis is drawn again by the reference (sinter_bridge.h:171-175), which shifts every
    later draw: the device scan detects it and hands over to the sequential scan."""
    from libgran_b200 import system as sysm

    scn = S.sintered_lattice()
    cfg = scn.cfg
    uv = unit_vectors_for(cfg, oracle, count=4096).copy()
    h_ij, _, _ = sysm.host_bond_table(scn.x0, cfg.r_part, cfg.critical_separation, uv)
    i, j = h_ij[4]  # third bond: make its vector the bond axis itself
    axis = scn.x0[j] - scn.x0[i]
    uv[2] = axis / np.linalg.norm(axis)
    h = sysm.host_bond_table(scn.x0, cfg.r_part, cfg.critical_separation, uv)
    d = sysm.device_bond_table(scn.x0, cfg.r_part, cfg.critical_separation, uv, device=0)
    for a, b in zip(h, d):
        assert np.array_equal(a, b)
