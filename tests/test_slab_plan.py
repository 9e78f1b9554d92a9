"""Host-side logic of the slab decomposition (libgran_b200/slab.py) on CPU tensors over gloo, world_size 2 and 3:
ownership is a partition, ghost coverage is complete for the Verlet radius, bond and contact-spring records travel
with their particles keyed by global id, the request-based halo lists move exactly the owners' values into the
ghost slots, and a second repartition after particles crossed a boundary (migration) keeps all of it."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from libgran_b200 import scenarios as S


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _check_tables(t, rank, world, bounds, gx, gstate, bonds, r_verlet, r_ghost, springs=None):
    from libgran_b200 import slab

    ij, conn, ln = bonds
    gid = t.gid.numpy()
    owned = t.owned.numpy()
    # local order is ascending global id, states are the global ones
    assert np.all(np.diff(gid) > 0)
    assert np.array_equal(t.state.numpy()[:, :12], gstate[gid][:, :12])
    # ownership is a partition and follows the bounds
    mine = torch.zeros(len(gx), dtype=torch.int64)
    mine[t.gid[t.owned]] = 1
    dist.all_reduce(mine)
    assert bool((mine == 1).all()), "ownership is not a partition"
    own_rank = slab.owner_of(torch.from_numpy(gx[gid][:, 0].copy()), bounds).numpy()
    assert np.all(own_rank[owned] == rank)
    assert np.array_equal(t.src.numpy(), own_rank)
    # ghost coverage: every particle within r_verlet of an owned particle is local
    local = np.zeros(len(gx), dtype=bool)
    local[gid] = True
    own_x = gx[gid[owned]]
    for k in range(0, len(own_x), 256):
        d = np.linalg.norm(own_x[k:k + 256, None, :] - gx[None, :, :], axis=2)
        need = (d < r_verlet).any(axis=0)
        assert np.all(local[need]), "a neighbor inside the Verlet radius is not local"
    # bonds of owned particles: same partners, connectors, lengths as the global table; mirrors resolve
    bptr, bj = t.bptr.numpy(), t.bj.numpy()
    order = np.lexsort((ij[:, 1], ij[:, 0]))
    ij_s, conn_s, ln_s = ij[order], conn[order], ln[order]
    starts = np.searchsorted(ij_s[:, 0], np.arange(len(gx) + 1))
    for li in np.nonzero(owned)[0]:
        g = gid[li]
        rows = slice(starts[g], starts[g + 1])
        loc = slice(bptr[li], bptr[li + 1])
        assert np.array_equal(gid[bj[loc]], ij_s[rows, 1])
        assert np.array_equal(t.bconn.numpy()[loc], conn_s[rows]) and np.array_equal(t.blen.numpy()[loc], ln_s[rows])
    rec_owner = np.repeat(np.arange(t.n), np.diff(bptr))
    brev = t.brev.numpy()
    assert np.array_equal(rec_owner[brev], bj) and np.array_equal(bj[brev], rec_owner)
    if springs is not None:  # {(gid_i, gid_j): payload[9]} for every ordered pair that had a record
        sg = t.sgid.numpy()
        seen = 0
        for li in np.nonzero(owned)[0]:
            for k in np.nonzero(sg[li] >= 0)[0]:
                key = (int(gid[li]), int(gid[sg[li, k]]))
                assert key in springs and np.array_equal(t.spr.numpy()[li, k], springs[key])
                seen += 1
        tot = torch.tensor([seen])
        dist.all_reduce(tot)
        assert int(tot) == len(springs)


def _halo_roundtrip(t, plan, rank):
    """What lg_halo_pack / NCCL / lg_halo_unpack do, in torch on CPU tensors: the ghosts must end up with the
    owners' values."""
    def f(g):  # a value only the owner 'knows'
        return torch.stack([g.double() * 3.0 + c for c in range(9)], dim=1)

    xvw = torch.full((t.n, 9), -1.0, dtype=torch.float64)
    xvw[t.owned] = f(t.gid[t.owned])
    conn = torch.full((t.n_bonds, 15), -1.0, dtype=torch.float64)
    rec_owner = torch.repeat_interleave(torch.arange(t.n), (t.bptr[1:] - t.bptr[:-1]).long())
    own_rec = t.owned[rec_owner]
    key = t.gid[rec_owner] * 1000003 + t.gid[t.bj.long()]
    conn[own_rec] = key[own_rec].double()[:, None] + torch.arange(15, dtype=torch.float64)[None, :]
    ops, bufs = [], {}
    for r in plan.peers:
        send = torch.cat([xvw[plan.send_idx[r].long()].reshape(-1), conn[plan.send_rec[r].long()].reshape(-1)])
        assert len(send) == plan.send_doubles(r)
        bufs[r] = torch.empty(plan.recv_doubles(r), dtype=torch.float64)
        ops += [dist.P2POp(dist.isend, send, r), dist.P2POp(dist.irecv, bufs[r], r)]
    if ops:
        for q in dist.batch_isend_irecv(ops):
            q.wait()
    for r in plan.peers:
        npart = len(plan.recv_idx[r])
        xvw[plan.recv_idx[r].long()] = bufs[r][:9 * npart].reshape(-1, 9)
        conn[plan.recv_rec[r].long()] = bufs[r][9 * npart:].reshape(-1, 15)
    assert torch.equal(xvw, f(t.gid)), "a ghost did not receive its owner's state"
    # connectors: a rank's own records, and the ghost records (j, i) an owned particle i reads as the mirror of (i, j);
    # records between two ghosts are never read (ghost rows are empty) and are not refreshed
    need = own_rec | t.owned[t.bj.long()]
    want = key.double()[:, None] + torch.arange(15, dtype=torch.float64)[None, :]
    assert torch.equal(conn[need], want[need]), "a mirror record did not receive its owner's connectors"


def _worker(rank, world, port):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from libgran_b200 import slab
        from libgran_b200.system import host_bond_table

        scn = S.multi_aggregate_collision(12)  # 1800 particles, 1800 bonds
        cfg = scn.cfg
        rng = np.random.Generator(np.random.MT19937(5))
        uv = rng.normal(size=(4 * cfg.n, 3))
        uv /= np.linalg.norm(uv, axis=1)[:, None]
        bonds = host_bond_table(scn.x0, cfg.r_part, cfg.critical_separation, uv)
        skin = 1.0 * cfg.r_part
        r_ghost = cfg.r_verlet + skin
        gstate = np.concatenate([scn.x0, scn.v0, scn.theta0, scn.omega0, np.zeros((cfg.n, 6))], axis=1)
        t0 = slab.block_tables(rank, world, scn.x0, scn.v0, scn.theta0, scn.omega0, 16, *bonds, extra_cols=10)  # one facet's wall history
        t1, bounds = slab.repartition_tables(t0, 0, r_ghost, world)
        assert len(bounds) == world - 1
        _check_tables(t1, rank, world, bounds, scn.x0, gstate, bonds, cfg.r_verlet, r_ghost)
        _halo_roundtrip(t1, slab.build_halo_plan(t1, rank, world), rank)

        # ---- migration: shift everything along x by a gid-dependent amount, give owned particles contact springs
        shift = (np.arange(cfg.n) % 7 - 3) * 0.4 * cfg.r_part + ((np.arange(cfg.n) // 150) % 2) * 45.0 * cfg.r_part
        gx2 = scn.x0.copy()
        gx2[:, 0] += shift
        t1.state[:, 0] = torch.from_numpy(gx2[t1.gid.numpy(), 0])
        gstate2 = gstate.copy()
        gstate2[:, 0] = gx2[:, 0]
        springs_local = {}
        own = torch.nonzero(t1.owned, as_tuple=True)[0].numpy()
        gidn = t1.gid.numpy()
        for li in own[::3]:
            d = np.linalg.norm(gx2[gidn] - gx2[gidn[li]], axis=1)
            near = np.argsort(d)[1:3]  # two nearest local particles (bonded neighbours: well inside the margin)
            for k, lj in enumerate(near):
                payload = np.arange(9, dtype=np.float64) + 100.0 * gidn[li] + gidn[lj]
                t1.sgid[li, k] = int(lj)
                t1.spr[li, k] = torch.from_numpy(payload)
                springs_local[(int(gidn[li]), int(gidn[lj]))] = payload
        gathered = [None] * world
        dist.all_gather_object(gathered, springs_local)
        springs = {k: v for d in gathered for k, v in d.items()}
        # connectors are particle-attached: move them with their particle so the global table stays the truth
        ij, conn, ln = bonds
        conn2 = conn.copy().reshape(-1, 5, 3)
        conn2[:, :, 0] += shift[ij[:, 0]][:, None]
        bonds2 = (ij, conn2.reshape(-1, 15), ln)
        rec_owner = torch.repeat_interleave(torch.arange(t1.n), (t1.bptr[1:] - t1.bptr[:-1]).long())
        t1.bconn.view(-1, 5, 3)[:, :, 0] += torch.from_numpy(shift[gidn[rec_owner.numpy()]])[:, None]
        # wall history of a facet (9 spring components + active flag per particle) rides in the state rows
        assert t1.state.shape[1] == 28 and bool((t1.state[:, 18:] == 0).all())
        wall = lambda g: g.double()[:, None] * 10.0 + torch.arange(10, dtype=torch.float64)[None, :]
        t1.state[:, 18:] = wall(t1.gid)
        t2, bounds2 = slab.repartition_tables(t1, 0, r_ghost, world)
        assert torch.equal(t2.state[:, 18:], wall(t2.gid)), "wall springs did not travel with their particles"
        moved = torch.tensor([int((t2.owned & (slab.owner_of(torch.from_numpy(scn.x0[t2.gid.numpy(), 0].copy()), bounds) != rank)).sum())])
        dist.all_reduce(moved)
        assert int(moved) > 0, "the test did not make any particle migrate"
        _check_tables(t2, rank, world, bounds2, gx2, gstate2, bonds2, cfg.r_verlet, r_ghost, springs)
        _halo_roundtrip(t2, slab.build_halo_plan(t2, rank, world), rank)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_plan_gloo(lib, world):
    mp.spawn(_worker, args=(world, _free_port()), nprocs=world, join=True)
