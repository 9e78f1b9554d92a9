"""Regenerates the fixtures under tests/golden/ from the reference (run in the build container only:
needs /root/reference and oracle/_ref/libref_oracle.so, i.e. `make -C oracle ref`).

    python tests/golden/make_golden.py

* ../../libgran_b200/data/aggregate_150.npy  the coordinate table of resources/aggregate.h:8-158 (units of r_part)
* <scenario>.npz         state of the UNMODIFIED reference (parity build, -ffp-contract=off) after
                         a number of lock-step steps: x, v, theta, omega, a, alpha, neighbor rows,
                         and the reference's own diagnostics (test/compute_energy.cpp,
                         test/mass_distribution.cpp) at the checkpoints.
"""
import os
import re
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("LIBGRAN_REFERENCE", "/root/reference")


def make_aggregate():
    txt = open(os.path.join(REF, "resources", "aggregate.h")).read()
    body = txt[txt.index("{") + 1: txt.rindex("}")]
    vals = [float(t) for t in re.findall(r"[-+]?\d+\.\d+(?:[eE][-+]?\d+)?", body)]
    arr = np.array(vals, dtype=np.float64).reshape(-1, 3)
    assert arr.shape == (150, 3), arr.shape
    np.save(os.path.join(os.path.dirname(os.path.dirname(HERE)), "libgran_b200", "data", "aggregate_150.npy"), arr)  # package data
    return arr


def run_reference(scn, checkpoints, facet_every_step=False):
    from oracle.bindings import RefSystem, ref_diagnostics

    cfg = scn.cfg
    ref = RefSystem(cfg, *scn.state)
    out = {}
    done = 0
    for cp in checkpoints:
        if facet_every_step:
            for s in range(done, cp):
                if cfg.use_neighbor_list and s % cfg.rebuild_every == 0:
                    ref.update_neighbor_list()
                ref.do_step(cfg.dt)
                ref.facet_update_positions(cfg.dt)
        else:
            # ref_run keeps its own step counter, so rebuilds stay on the global % rebuild_every grid
            ref.run(cfg.dt, cp - done, cfg.rebuild_every if cfg.use_neighbor_list else 0)
        done = cp
        for name in ("x", "v", "theta", "omega", "a", "alpha"):
            out[f"{name}_{cp}"] = ref.get(name)
        d = ref_diagnostics(ref.get("x"), ref.get("v"), ref.get("omega"), cfg.mass, cfg.inertia)
        out[f"diag_{cp}"] = np.array([d["ke"], d["p"], d["l"], d["rg"], *d["com"]])
    row_ptr, cols = ref.rows()
    out["row_ptr"], out["cols"] = row_ptr, cols
    out["checkpoints"] = np.array(checkpoints)
    ref.close()
    return out


def main():
    make_aggregate()
    from libgran_b200 import scenarios as S

    cases = {
        "slab_contact_allpairs": (S.slab_and_projectile(S.MODEL_CONTACT, neighbor_list=False), [1, 2, 600, 1500]),
        "slab_hamaker_neighbor": (S.slab_and_projectile(S.MODEL_CONTACT | S.MODEL_HAMAKER, neighbor_list=True), [1, 2, 600, 1500]),
        "sintered_lattice": (S.sintered_lattice(), [1, 2, 500, 1000]),
        "aggregate_sinter_hamaker": (S.aggregate(), [1, 2, 500, 1000]),
        "surface_drop": (S.surface_drop(), [1, 2, 40000, 60000]),
    }
    # branch-coverage inputs (contacts forming / sliding / separating, facet plane + edge + vertex, moving facet)
    stress = {
        "stress_contact_hamaker_facet": S.stress_block(models=S.MODEL_CONTACT | S.MODEL_HAMAKER | S.MODEL_FACET),
        "stress_sinter_hamaker_facet": S.stress_block(models=S.MODEL_SINTER | S.MODEL_HAMAKER | S.MODEL_FACET),
        "stress_sinter_allpairs": S.stress_block(nx=4, ny=4, nz=3, models=S.MODEL_SINTER, neighbor_list=False),
    }
    for name, scn in stress.items():
        out = run_reference(scn, [1, 2, 50, 200], facet_every_step=True)
        out["x0"], out["v0"], out["omega0"] = scn.x0, scn.v0, scn.omega0
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "ok")
    for name, (scn, cps) in cases.items():
        out = run_reference(scn, cps)
        out["x0"], out["v0"] = scn.x0, scn.v0
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "ok", {k: v.shape for k, v in out.items() if k.startswith("x_")})


if __name__ == "__main__":
    main()
