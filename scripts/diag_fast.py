import sys, numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from libgran_b200 import scenarios as S
from libgran_b200.system import from_config
from oracle import bindings as ob
from helpers import unit_vectors_for
scn = S.stress_block(14, 14, 10, models=S.MODEL_SINTER | S.MODEL_HAMAKER | S.MODEL_FACET)
cfg = scn.cfg
uv = unit_vectors_for(cfg, ob)
for moves in (False, True):
    e = from_config(cfg, *scn.state, unit_vectors=uv, device=0, math_mode=0)
    f = from_config(cfg, *scn.state, unit_vectors=uv, device=0, math_mode=1)
    print("moves", moves)
    for k in range(1, 13):
        for g in (e, f):
            g.run(cfg.dt, 1, cfg.rebuild_every, moves)
        xe, xf, ve, vf = e.get("x"), f.get("x"), e.get("v"), f.get("v")
        we, wf = e.get("omega"), f.get("omega")
        dv = np.abs(vf - ve).max() / np.abs(ve).max(); dw = np.abs(wf - we).max() / np.abs(we).max()
        i = np.unravel_index(np.abs(vf - ve).argmax(), ve.shape)
        print(k, "dv %.2e dw %.2e" % (dv, dw), "worst particle", i[0], "contacts", e.counters()["contacts"], f.counters()["contacts"])
