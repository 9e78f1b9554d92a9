"""2-rank probe: first step at which a slab run with a moving facet differs from one GPU (diagnostic, not a test)"""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from libgran_b200 import scenarios as S
STEPS = 30
CHUNK = int(os.environ.get("PROBE_CHUNK", "1"))


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
    for k in range(0, STEPS, CHUNK):
        s.run(cfg.dt, CHUNK, cfg.rebuild_every, True)
        res[f"x_{k + CHUNK}"] = s.get("x"); res[f"v_{k + CHUNK}"] = s.get("v")
        verts = np.zeros(9); s._check(s.lib.lg_facet_get_vertices(s.h, 0, verts.ctypes.data_as(__import__("ctypes").POINTER(__import__("ctypes").c_double))))
        res[f"f_{k + CHUNK}"] = verts
    own = np.zeros(cfg.n, dtype=np.int64); own[s.gid[s.own_idx].cpu().numpy()] = rank + 1
    t = torch.from_numpy(own).to(s.device); dist.all_reduce(t)
    if rank == 0:
        np.savez(out, owner=t.cpu().numpy(), reps=s.repartitions, **res)
    s.close(); dist.destroy_process_group()


if __name__ == "__main__":
    import torch.multiprocessing as mp
    out = "/tmp/probe2.npz"
    mp.spawn(worker, args=(2, 29578, out), nprocs=2, join=True)
    from libgran_b200.system import from_config
    scn = scenario(); cfg = scn.cfg
    g = from_config(cfg, *scn.state, device=0)
    got = np.load(out)
    print("repartitions", got["reps"], "chunk", CHUNK)
    zf = float(np.asarray(cfg.facet_vertices)[0, 2])
    for k in range(0, STEPS, CHUNK):
        g.run(cfg.dt, CHUNK, cfg.rebuild_every, True)
        x, v, fv = g.get("x"), g.get("v"), g.facet_vertices(0).reshape(9)
        kk = k + CHUNK
        bx = np.argwhere((got[f"x_{kk}"] != x).any(axis=1)).ravel(); bv = np.argwhere((got[f"v_{kk}"] != v).any(axis=1)).ravel()
        print("step", kk, "x differs", len(bx), "v differs", len(bv), "facet equal", np.array_equal(got[f"f_{kk}"], fv),
              "max dv", np.abs(got[f"v_{kk}"] - v).max())
        if len(bv):
            b = bv[:6]
            print("   ids", b, "owner", got["owner"][b], "x/r", np.round(scn.x0[b, 0] / S.R_PART, 2), "(z - z_facet)/r", np.round((scn.x0[b, 2] - zf) / S.R_PART, 2))
            break
