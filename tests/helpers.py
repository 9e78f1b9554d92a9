"""Shared helpers of the parity tests: drive two systems in lock-step and compare them bit for bit."""
import os

import numpy as np

from libgran_b200 import scenarios as S

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIELDS = ("x", "v", "theta", "omega", "a", "alpha")


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def golden_scenarios():
    """name -> (scenario, moves facet every step) for every committed .npz (tests/golden/make_golden.py)."""
    return {
        "slab_contact_allpairs": (S.slab_and_projectile(S.MODEL_CONTACT, neighbor_list=False), False),
        "slab_hamaker_neighbor": (S.slab_and_projectile(S.MODEL_CONTACT | S.MODEL_HAMAKER, neighbor_list=True), False),
        "sintered_lattice": (S.sintered_lattice(), False),
        "aggregate_sinter_hamaker": (S.aggregate(), False),
        "surface_drop": (S.surface_drop(), False),
        "stress_contact_hamaker_facet": (S.stress_block(models=S.MODEL_CONTACT | S.MODEL_HAMAKER | S.MODEL_FACET), True),
        "stress_sinter_hamaker_facet": (S.stress_block(models=S.MODEL_SINTER | S.MODEL_HAMAKER | S.MODEL_FACET), True),
        "stress_sinter_allpairs": (S.stress_block(nx=4, ny=4, nz=3, models=S.MODEL_SINTER, neighbor_list=False), True),
    }


def advance(sysm, cfg, start, stop, move_facets):
    """Driver loop of test/hamaker_neighbor.cpp:120-124 (+ facet.update_positions as a driver would call it)."""
    every = cfg.rebuild_every if cfg.use_neighbor_list else 0
    if hasattr(sysm, "lib") and hasattr(sysm.lib, "lg_run"):
        sysm.run(cfg.dt, stop - start, every, move_facets)
        return
    if not move_facets:
        sysm.run(cfg.dt, stop - start, every)
        return
    for s in range(start, stop):
        if every and s % every == 0:
            sysm.update_neighbor_list()
        sysm.do_step(cfg.dt)
        sysm.facet_update_positions(cfg.dt)


def assert_fields_equal(a, b, what, exact=True, rtol=0.0):
    for f in FIELDS:
        fa, fb = a.get(f), b.get(f)
        if exact:
            if not np.array_equal(fa, fb):
                bad = np.argwhere(fa != fb)
                scale = max(np.abs(fb).max(), 1e-300)
                raise AssertionError(
                    f"{what}: field {f} differs at {len(bad)} entries, first {bad[0]}, "
                    f"max rel {np.abs(fa - fb).max() / scale:.3e}")
        else:
            scale = max(np.abs(fb).max(), 1e-300)
            err = np.abs(fa - fb).max() / scale
            assert err <= rtol, f"{what}: field {f} rel err {err:.3e} > {rtol:.1e}"


def unit_vectors_for(cfg, oracle_bindings, count=None):
    """The unit vectors the reference's generator (mt19937_64(seed), test/sintered.cpp:29-39) hands to the
    sinter functor, so that the CUDA path builds the same connectors as the oracle."""
    if not (cfg.model_mask & S.MODEL_SINTER):
        return None
    return oracle_bindings.oracle_unit_vectors(cfg.rng_seed, count or max(16 * cfg.n, 1024))
