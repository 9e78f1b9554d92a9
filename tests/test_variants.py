"""SURVEY.md section 8f variants.

CPU (-m "not gpu"): the oracle restatement of alt_sinter_functor, limited_range / simplified Hamaker, forward Euler and
the fixed-particle step handler is bit-identical to the UNMODIFIED reference (oracle/_ref) in lock-step.
GPU (-m gpu): the CUDA path through the C ABI is bit-identical to the oracle on the same cases, including the hidden
per-bond springs of alt_sinter_functor and the bond SET (cycle prevention)."""
import numpy as np
import pytest

from libgran_b200 import scenarios as S

from helpers import FIELDS, advance, assert_fields_equal
from variant_cases import CASES


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_variants_match_live_reference(oracle, case):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference: make -C oracle ref)")
    _, make, steps, ref_ok = case
    if not ref_ok:
        pytest.skip("combination not instantiated in oracle/ref_harness.cpp (fixed-particle handler: tutorial text, not a reference header)")
    scn = make()
    r = oracle.RefSystem(scn.cfg, *scn.state)
    o = oracle.OracleSystem(scn.cfg, *scn.state)
    chunk = max(1, steps // 5)
    for s in range(0, steps, chunk):
        stop = min(steps, s + chunk)
        advance(r, scn.cfg, s, stop, False)
        advance(o, scn.cfg, s, stop, False)
        for f in FIELDS:
            assert np.array_equal(r.get(f), o.get(f)), f"{scn.name}: {f} at step {stop}"
    n = scn.cfg.n
    for i in range(min(n, 40)):
        for j in range(min(n, 40)):
            if i != j:
                assert np.array_equal(r.contact_springs(i, j), o.contact_springs(i, j)), f"springs ({i},{j})"
                assert (r.bond(i, j) is None) == (o.bond(i, j) is None), f"bond flag ({i},{j})"


def test_oracle_fixed_particle_semantics(oracle):
    """Custom-step-handler.md:95-103: increment_v of the fixed particle is dropped -- its velocity never changes (also not
    by the first-step half kick), while it still moves with that velocity; everything else is the plain run with the
    fixed particle's velocity history replaced."""
    from variant_cases import fixed

    scn = fixed(S.MODEL_SINTER | S.MODEL_HAMAKER)
    o = oracle.OracleSystem(scn.cfg, *scn.state)
    o.run(scn.cfg.dt, 50, scn.cfg.rebuild_every)
    k = scn.cfg.fixed_particle
    assert np.array_equal(o.get("v")[k], scn.v0[k])
    assert not np.array_equal(o.get("x")[k], scn.x0[k])
    free = oracle.OracleSystem(scn.cfg.with_(fixed_particle=-1), *scn.state)
    free.run(scn.cfg.dt, 50, scn.cfg.rebuild_every)
    assert not np.array_equal(free.get("v")[k], scn.v0[k])


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_variants_match_oracle(lib, oracle, case):
    from libgran_b200.system import from_config
    from helpers import unit_vectors_for

    _, make, steps, _ = case
    scn = make()
    gpu = from_config(scn.cfg, *scn.state, unit_vectors=unit_vectors_for(scn.cfg, oracle), device=0)
    cpu = oracle.OracleSystem(scn.cfg, *scn.state)
    chunk = max(1, steps // 5)
    for s in range(0, steps, chunk):
        stop = min(steps, s + chunk)
        advance(gpu, scn.cfg, s, stop, False)
        advance(cpu, scn.cfg, s, stop, False)
        assert_fields_equal(gpu, cpu, f"{scn.name} step {stop}")
    if scn.cfg.use_neighbor_list:
        (gp, gc), (cp, cc) = gpu.rows(), cpu.rows()
        assert np.array_equal(gp, cp) and np.array_equal(gc, cc)
    assert gpu.counters()["bonds"] == cpu.n_bonds(), "bond set size (alt: after cycle prevention)"
    n = scn.cfg.n
    alt = bool(scn.cfg.model_mask & S.MODEL_ALT_SINTER)
    for i in range(min(n, 30)):
        for j in range(min(n, 30)):
            if i == j:
                continue
            bonded = cpu.bond(i, j) is not None
            if alt:
                got = gpu.alt_bond(i, j)
                assert (got is not None) == bonded, f"alt bond flag ({i},{j})"
                if bonded:
                    assert np.array_equal(got, cpu.contact_springs(i, j)), f"alt bond springs ({i},{j})"
                    continue
            assert np.array_equal(gpu.contact_springs(i, j), cpu.contact_springs(i, j)), f"springs ({i},{j})"


@pytest.mark.gpu
def test_gpu_central_attraction(lib):
    """central_attraction_functor (Writerside/topics/Custom-unary-force-model.md:99-100): a_i = k_a / mass * (x_i - x0),
    with x0 the index-order mean the tutorial's update_center_of_mass computes; checked against the same expression in numpy."""
    from libgran_b200.system import from_config

    scn = S.slab_and_projectile(S.MODEL_CONTACT | S.MODEL_HAMAKER, True)
    plain = from_config(scn.cfg, *scn.state, device=0, variant=0)
    cen = from_config(scn.cfg, *scn.state, device=0)
    k_a = -8.0e-6
    x0 = np.zeros(3)
    for p in scn.x0:  # serial sum in index order, then / n
        x0 = x0 + p
    x0 = x0 / float(len(scn.x0))
    cen.set_central_attraction(k_a, x0)
    for s in (plain, cen):
        s.update_neighbor_list()
        s.do_step(0.0)
    want = plain.get("a") + (k_a / scn.cfg.mass) * (scn.x0 - x0)
    assert np.array_equal(cen.get("a"), want)
    assert np.array_equal(cen.get("alpha"), plain.get("alpha"))
    assert cen.last_step_kernel() == "force_simple_kernel"


@pytest.mark.gpu
def test_gpu_variant_guards(lib):
    from libgran_b200.system import LgError, from_config

    scn = S.slab_and_projectile(S.MODEL_CONTACT, True)
    g = from_config(scn.cfg.with_(integrator=1), *scn.state, device=0)
    with pytest.raises(LgError):
        g.set_kernel_variant(1)  # the fused kernels implement the half-step Verlet schedule only
    with pytest.raises(LgError):
        g.set_math_mode(1)
    g.do_step(scn.cfg.dt)
    with pytest.raises(LgError):
        g.set_integrator(0)  # a constructor choice
