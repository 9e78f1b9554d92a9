"""Set-ups of the SURVEY.md section 8f variants shared by the CPU (oracle vs the live reference) and GPU (CUDA path vs
oracle) tests: alt_sinter_functor, the Hamaker variants, rotational_forward_euler, fixed_particle_step_handler."""
import numpy as np

from libgran_b200 import scenarios as S


def _kick(scn, seed=7, scale=2.0):
    rng = np.random.default_rng(seed)
    scn.v0 = np.ascontiguousarray(rng.uniform(-scale, scale, size=scn.x0.shape))
    scn.omega0 = np.ascontiguousarray(rng.uniform(-1e7, 1e7, size=scn.x0.shape))
    return scn


def alt_lattice(neighbor_list, hamaker):
    return S.alt_sintered_lattice(neighbor_list=neighbor_list, hamaker=hamaker)


def alt_block(hamaker=False):
    """A jittered 6x6x5 block bonded by alt_sinter_functor: many candidate bonds close cycles (cycle prevention matters),
    the non-bonded touching pairs run the embedded contact model."""
    base = S.stress_block(models=S.MODEL_CONTACT | S.MODEL_HAMAKER, neighbor_list=True)
    alt = S.alt_sintered_lattice().cfg
    mask = S.MODEL_ALT_SINTER | (S.MODEL_HAMAKER if hamaker else 0)
    cfg = base.cfg.with_(model_mask=mask, alt_bond=alt.alt_bond, critical_separation=0.08 * S.R_PART)
    return S.Scenario("alt_block", cfg, base.x0, base.v0, base.theta0, base.omega0)


def hamaker_variant(kind, neighbor_list=True):
    scn = S.slab_and_projectile(S.MODEL_CONTACT | S.MODEL_HAMAKER, neighbor_list)
    scn.cfg = scn.cfg.with_(hamaker_kind=kind, hamaker_cutoff=3.3 * S.R_PART)
    return scn


def euler(models, neighbor_list=True):
    scn = S.sintered_lattice(models, neighbor_list) if models & S.MODEL_SINTER else S.slab_and_projectile(models, neighbor_list)
    scn.cfg = scn.cfg.with_(integrator=1)
    return _kick(scn)


def fixed(models, integrator=0):
    scn = S.sintered_lattice(models, True) if models & S.MODEL_SINTER else S.slab_and_projectile(models, True)
    scn.cfg = scn.cfg.with_(fixed_particle=3, integrator=integrator)
    return _kick(scn)


#: (id, factory, steps, the reference harness can run it)
CASES = [
    ("alt lattice allpairs", lambda: alt_lattice(False, False), 400, True),
    ("alt lattice nl", lambda: alt_lattice(True, False), 400, True),
    ("alt lattice nl + hamaker", lambda: alt_lattice(True, True), 400, False),
    ("alt block nl", alt_block, 150, True),
    ("alt block nl + hamaker", lambda: alt_block(True), 150, False),
    ("hamaker limited range nl", lambda: hamaker_variant(1), 300, True),
    ("hamaker simplified nl", lambda: hamaker_variant(2), 300, True),
    ("hamaker simplified allpairs", lambda: hamaker_variant(2, False), 200, True),
    ("euler contact+hamaker nl", lambda: euler(S.MODEL_CONTACT | S.MODEL_HAMAKER), 300, True),
    ("euler sinter+hamaker allpairs", lambda: euler(S.MODEL_SINTER | S.MODEL_HAMAKER, False), 300, True),
    ("fixed particle sinter+hamaker", lambda: fixed(S.MODEL_SINTER | S.MODEL_HAMAKER), 300, False),
    ("fixed particle euler contact", lambda: fixed(S.MODEL_CONTACT, 1), 300, False),
]
