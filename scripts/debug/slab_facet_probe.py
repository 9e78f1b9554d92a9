"""2-rank probe: where does a slab run with a facet differ from one GPU?  (diagnostic, not a test)"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from libgran_b200 import scenarios as S


def scenario():
    return S.stress_block(14, 14, 10, models=S.MODEL_CONTACT | S.MODEL_HAMAKER | S.MODEL_FACET)


def worker(rank, world, port, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    from libgran_b200.slab import SlabSystem
    scn = scenario(); cfg = scn.cfg
    s = SlabSystem(cfg, *scn.state, device=rank, math_mode=0)
    res = {}
    s.update_neighbor_list()
    s._check(s.lib.lg_first_half_kick(s.h, cfg.dt)); s.halo_exchange(); s.kicked = True
    res["v_half"] = s.get("v"); res["a0"] = s.get("a")
    s._check(s.lib.lg_do_step(s.h, cfg.dt)); s.halo_exchange(); s.steps += 1
    for f in ("x", "v", "omega", "a", "alpha"):
        res[f + "_1"] = s.get(f)
    own = np.zeros(cfg.n, dtype=np.int64); g = s.gid[s.own_idx].cpu().numpy()
    own[g] = 1
    t = torch.from_numpy(own).to(s.device) * (rank + 1)
    dist.all_reduce(t)
    if rank == 0:
        np.savez(out, owner=t.cpu().numpy(), bounds=np.asarray(s.bounds.cpu() if hasattr(s.bounds, "cpu") else s.bounds), **res)
    s.close(); dist.destroy_process_group()


if __name__ == "__main__":
    import torch.multiprocessing as mp
    out = "/tmp/probe.npz"
    mp.spawn(worker, args=(2, 29577, out), nprocs=2, join=True)
    from libgran_b200.system import from_config
    scn = scenario(); cfg = scn.cfg
    g = from_config(cfg, *scn.state, device=0)
    g.update_neighbor_list()
    g._check(g.lib.lg_first_half_kick(g.h, cfg.dt))
    ref = {"v_half": g.get("v"), "a0": g.get("a")}
    g.do_step(cfg.dt)
    for f in ("x", "v", "omega", "a", "alpha"):
        ref[f + "_1"] = g.get(f)
    got = np.load(out)
    print("bounds", got["bounds"], "x range", scn.x0[:, 0].min(), scn.x0[:, 0].max())
    zf = float(np.asarray(cfg.facet_vertices)[0, 2])
    for k, v in ref.items():
        bad = np.argwhere((got[k] != v).any(axis=1)).ravel()
        print(k, "differs at", len(bad), "particles", "max abs", np.abs(got[k] - v).max() if len(bad) else 0.0)
        if len(bad):
            b = bad[:8]
            print("   ids", b, "owner", got["owner"][b], "x", scn.x0[b, 0] / S.R_PART, "z - z_facet (r)", (scn.x0[b, 2] - zf) / S.R_PART)
