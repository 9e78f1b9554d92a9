"""1-D slab decomposition of a granular system over several GPUs.

One process per GPU.  ``torch.distributed`` is the plumbing (NCCL between GPUs, gloo in the CPU tests); every
per-particle computation stays in the CUDA library behind the C ABI.  The reference has no distributed path, so
nothing here restates reference code; what must hold is that a decomposed run returns the *same bits* as the
single-GPU run, which it does because each rank keeps its local particles ordered by global id (ascending local
neighbor index == ascending global index == the reference's accumulation order).

Layout of this module
  * plan functions (pure torch, any device, any backend): slab bounds, owners, ghost destinations, the record
    streams that move particles between ranks, re-localisation of what arrived, halo request lists.  They are
    what tests/test_slab_plan.py drives on CPU tensors with gloo, world_size 2.
  * ``SlabSystem``: the GPU driver -- owns one ``lg_system`` per rank, repartitions when particles have moved more
    than half the skin, exchanges ghost x, v, omega and bond connectors after every step.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch
import torch.distributed as dist

# columns of the three record streams (all float64 rows; integer columns are int64 bit patterns)
P_INT, P_DBL = 3, 18   # gid, kind (0 owned / 1 ghost), new owner rank | x v theta omega a alpha
B_INT, B_DBL = 2, 20   # gid_i, gid_j | 5 connectors of i (15), 5 rest lengths
C_INT, C_DBL = 2, 9    # gid_i, gid_j | 3 history springs of the ordered pair (i, j)


# =================================================================================================
# small helpers
# =================================================================================================
def _i2f(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.int64).contiguous().view(torch.float64)


def _f2i(t: torch.Tensor) -> torch.Tensor:
    return t.contiguous().view(torch.int64)


def ragged_expand(starts: torch.Tensor, counts: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """For segments [starts[k], starts[k] + counts[k]) return (segment id, element index) of every element."""
    counts = counts.to(torch.int64)
    total = int(counts.sum())
    dev = counts.device
    if total == 0:
        z = torch.zeros(0, dtype=torch.int64, device=dev)
        return z, z.clone()
    seg = torch.repeat_interleave(torch.arange(len(counts), device=dev), counts)
    first = torch.cumsum(counts, 0) - counts
    offs = torch.arange(total, device=dev) - first[seg]
    return seg, starts.to(torch.int64)[seg] + offs


def lookup_sorted(sorted_keys: torch.Tensor, query: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Positions of `query` in ascending `sorted_keys` and a mask of the ones actually present."""
    if len(sorted_keys) == 0:
        return torch.zeros_like(query), torch.zeros_like(query, dtype=torch.bool)
    pos = torch.searchsorted(sorted_keys, query)
    pos_c = pos.clamp(max=len(sorted_keys) - 1)
    return pos_c, sorted_keys[pos_c] == query


# =================================================================================================
# plan: who owns what, who needs which ghosts
# =================================================================================================
def slab_bounds(coord_owned: torch.Tensor, world: int, group=None) -> torch.Tensor:
    """Inner slab boundaries (world - 1 values) at particle-count quantiles of the slab-axis coordinate of ALL
    particles.  Every rank passes the coordinates of the particles it currently owns and gets identical bounds."""
    dev = coord_owned.device
    if world == 1:
        return torch.zeros(0, dtype=torch.float64, device=dev)
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    mine = torch.tensor([len(coord_owned)], dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(counts, mine, group=group)
    cmax = int(counts.max())
    padded = torch.full((cmax,), float("inf"), dtype=torch.float64, device=dev)
    padded[: len(coord_owned)] = coord_owned
    everything = torch.empty(world * cmax, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(everything, padded, group=group)
    xs, _ = torch.sort(everything)  # the +inf padding sorts to the end
    n = int(counts.sum())
    cuts = [(k * n) // world for k in range(1, world)]
    idx = torch.tensor(cuts, dtype=torch.int64, device=dev).clamp(1, max(n - 1, 1))
    return 0.5 * (xs[idx - 1] + xs[idx])


def owner_of(coord: torch.Tensor, inner_bounds: torch.Tensor) -> torch.Tensor:
    """Rank owning each coordinate: the number of inner bounds <= coord."""
    if len(inner_bounds) == 0:
        return torch.zeros(len(coord), dtype=torch.int64, device=coord.device)
    return torch.bucketize(coord, inner_bounds, right=True)


def ghost_destinations(coord: torch.Tensor, owner: torch.Tensor, inner_bounds: torch.Tensor, r_ghost: float,
                       world: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """(particle index, rank) for every rank other than the owner whose slab, widened by r_ghost on both sides,
    contains the particle."""
    dev = coord.device
    if world == 1 or len(coord) == 0:
        z = torch.zeros(0, dtype=torch.int64, device=dev)
        return z, z.clone()
    ninf = torch.tensor([-float("inf")], dtype=torch.float64, device=dev)
    pinf = torch.tensor([float("inf")], dtype=torch.float64, device=dev)
    lo = torch.cat([ninf, inner_bounds]) - r_ghost   # [world]
    hi = torch.cat([inner_bounds, pinf]) + r_ghost
    inside = (coord[:, None] >= lo[None, :]) & (coord[:, None] < hi[None, :])
    inside[torch.arange(len(coord), device=dev), owner] = False
    p, r = torch.nonzero(inside, as_tuple=True)
    return p, r


def all_to_all_rows(rows: torch.Tensor, dest: torch.Tensor, world: int, group=None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Send row k of `rows` (float64 [m, c]) to rank dest[k]; returns the rows received (grouped by source rank in
    rank order, each group in the sender's order) and the source rank of every received row."""
    dev = rows.device
    order = torch.argsort(dest, stable=True)
    rows = rows[order].contiguous()
    send_counts = torch.bincount(dest, minlength=world).to(torch.int64)
    recv_counts = torch.empty_like(send_counts)
    if world == 1:
        return rows, torch.zeros(len(rows), dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv_counts, send_counts, group=group)
    sc, rc = send_counts.tolist(), recv_counts.tolist()
    out = torch.empty((sum(rc), rows.shape[1]), dtype=rows.dtype, device=dev)
    dist.all_to_all_single(out, rows, rc, sc, group=group)
    src = torch.repeat_interleave(torch.arange(world, device=dev), recv_counts)
    return out, src


# =================================================================================================
# local tables and the record streams that move them
# =================================================================================================
@dataclass
class LocalTables:
    """Everything one rank holds, in local order (ascending global id).  Pure data, any device."""

    gid: torch.Tensor      # int64 [n]
    owned: torch.Tensor    # bool  [n]
    src: torch.Tensor      # int64 [n]  rank the record came from (the owner, for ghosts)
    state: torch.Tensor    # float64 [n, 18 + 10 F]  x v theta omega a alpha, then per facet: 9 wall-spring components + active flag
    S: int
    sgid: torch.Tensor     # int32 [n, S]  local partner index or -1
    spr: torch.Tensor      # float64 [n, S, 9]
    bptr: torch.Tensor     # int32 [n + 1]
    bj: torch.Tensor       # int32 [nb]  local partner index, ascending inside a row
    brev: torch.Tensor     # int32 [nb]  record index of the mirror (j, i)
    bconn: torch.Tensor    # float64 [nb, 15]
    blen: torch.Tensor     # float64 [nb, 5]

    @property
    def n(self) -> int:
        return len(self.gid)

    @property
    def n_bonds(self) -> int:
        return len(self.bj)


def bond_rows(t: LocalTables) -> Tuple[torch.Tensor, torch.Tensor]:
    """(row pointer int64 [n + 1], bond-stream rows [nb, 22] in local record order) of a rank's bond table."""
    dev = t.gid.device
    if hasattr(t, "global_bonds"):  # initial block distribution: records already carry global ids
        B = t.global_bonds
        li = _f2i(B[:, 0].contiguous()) - int(t.gid[0]) if len(B) else torch.zeros(0, dtype=torch.int64, device=dev)
        o = torch.argsort(li, stable=True)
        bptr = torch.zeros(t.n + 1, dtype=torch.int64, device=dev)
        if len(li):
            bptr[1:] = torch.cumsum(torch.bincount(li, minlength=t.n), 0)
        return bptr, B[o]
    bptr = t.bptr.to(torch.int64)
    if t.n_bonds == 0:
        return bptr, torch.zeros((0, B_INT + B_DBL), dtype=torch.float64, device=dev)
    rec_owner = torch.repeat_interleave(torch.arange(t.n, device=dev), bptr[1:] - bptr[:-1])
    rows = torch.cat([_i2f(t.gid[rec_owner])[:, None], _i2f(t.gid[t.bj.to(torch.int64)])[:, None], t.bconn, t.blen], dim=1)
    return bptr, rows


def make_streams(t: LocalTables, coord_axis: int, inner_bounds: torch.Tensor, r_ghost: float, world: int):
    """Record streams for a repartition: every OWNED particle goes to its new owner (full record with its contact
    springs and bond rows) and, as a ghost (state + bond rows), to every other rank whose widened slab holds it.
    Returns [(rows, dest)] for the particle, bond and contact streams."""
    dev = t.gid.device
    own_idx = torch.nonzero(t.owned, as_tuple=True)[0]
    coord = t.state[own_idx, coord_axis]
    owner = owner_of(coord, inner_bounds)
    gp, gr = ghost_destinations(coord, owner, inner_bounds, r_ghost, world)
    ship_p = torch.cat([own_idx, own_idx[gp]])           # local index of the shipped particle
    ship_dest = torch.cat([owner, gr])
    ship_kind = torch.cat([torch.zeros_like(owner), torch.ones_like(gr)])
    ship_owner = torch.cat([owner, owner[gp]])
    # particle stream
    P = torch.cat([_i2f(t.gid[ship_p])[:, None], _i2f(ship_kind)[:, None], _i2f(ship_owner)[:, None], t.state[ship_p]], dim=1)
    # bond stream: all records of every shipped particle
    bptr, rows = bond_rows(t)
    nb_of = bptr[1:] - bptr[:-1]
    seg, rec = ragged_expand(bptr[ship_p], nb_of[ship_p])
    B = rows[rec] if len(rec) else torch.zeros((0, B_INT + B_DBL), dtype=torch.float64, device=dev)
    B_dest = ship_dest[seg]
    # contact stream: active slots of owned particles, to the new owner only
    sg = t.sgid[own_idx].to(torch.int64)                  # [n_own, S]
    act_p, act_k = torch.nonzero(sg >= 0, as_tuple=True)
    ci = t.gid[own_idx[act_p]]
    cj = t.gid[sg[act_p, act_k]]
    Cs = torch.cat([_i2f(ci)[:, None], _i2f(cj)[:, None], t.spr[own_idx[act_p], act_k]], dim=1) if len(act_p) else \
        torch.zeros((0, C_INT + C_DBL), dtype=torch.float64, device=dev)
    C_dest = owner[act_p]
    return (P, ship_dest), (B, B_dest), (Cs, C_dest)


def relocalize(P: torch.Tensor, P_src: torch.Tensor, B: torch.Tensor, Cs: torch.Tensor, S: int) -> LocalTables:
    """Build a rank's local tables from the records it received (its own included)."""
    dev = P.device
    gid_all = _f2i(P[:, 0].contiguous())
    order = torch.argsort(gid_all, stable=True)
    gid = gid_all[order]
    if len(gid) > 1 and bool((gid[1:] == gid[:-1]).any()):
        raise RuntimeError("slab repartition: a particle arrived twice on one rank")
    kind = _f2i(P[:, 1].contiguous())[order]
    n = len(gid)
    owned = kind == 0
    state = P[order, P_INT:].contiguous()
    src = _f2i(P[:, 2].contiguous())[order]  # the particle's (new) owner rank
    # ---- bonds ---------------------------------------------------------------------------------
    bi_g, bj_g = _f2i(B[:, 0].contiguous()), _f2i(B[:, 1].contiguous())
    li, ok_i = lookup_sorted(gid, bi_g)
    lj, ok_j = lookup_sorted(gid, bj_g)
    if len(li) and not bool(ok_i.all()):
        raise RuntimeError("slab repartition: bond record without its particle")
    if len(li) and bool((owned[li] & ~ok_j).any()):
        raise RuntimeError("slab repartition: a bonded partner of an owned particle is outside the ghost margin "
                           "(increase skin)")
    keep = ok_j
    li, lj, Bk = li[keep], lj[keep], B[keep]
    key = li * max(n, 1) + lj
    bo = torch.argsort(key, stable=True)
    key, li, lj, Bk = key[bo], li[bo], lj[bo], Bk[bo]
    if len(key) > 1 and bool((key[1:] == key[:-1]).any()):
        raise RuntimeError("slab repartition: duplicate bond record")
    bptr = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    if len(li):
        bptr[1:] = torch.cumsum(torch.bincount(li, minlength=n), 0)
    rev_pos, rev_ok = lookup_sorted(key, lj * max(n, 1) + li)
    if len(key) and not bool(rev_ok.all()):
        raise RuntimeError("slab repartition: bond record without its mirror")
    # ---- contact springs ---------------------------------------------------------------------------
    sgid = torch.full((n, S), -1, dtype=torch.int32, device=dev)
    spr = torch.zeros((n, S, 9), dtype=torch.float64, device=dev)
    if len(Cs):
        ci, ok_ci = lookup_sorted(gid, _f2i(Cs[:, 0].contiguous()))
        cj, ok_cj = lookup_sorted(gid, _f2i(Cs[:, 1].contiguous()))
        if not bool((ok_ci & ok_cj).all()):
            raise RuntimeError("slab repartition: a contact partner is outside the ghost margin")
        co = torch.argsort(ci * max(n, 1) + cj, stable=True)
        ci, cj, Ck = ci[co], cj[co], Cs[co]
        first = torch.cumsum(torch.bincount(ci, minlength=n), 0) - torch.bincount(ci, minlength=n)
        slot = torch.arange(len(ci), device=dev) - first[ci]
        if int(slot.max()) >= S:
            raise RuntimeError("slab repartition: more simultaneous contacts than max_contacts")
        sgid[ci, slot] = cj.to(torch.int32)
        spr[ci, slot] = Ck[:, C_INT:]
    return LocalTables(gid=gid, owned=owned, src=src, state=state, S=S, sgid=sgid, spr=spr,
                       bptr=bptr.to(torch.int32), bj=lj.to(torch.int32), brev=rev_pos.to(torch.int32),
                       bconn=Bk[:, B_INT:B_INT + 15].contiguous(), blen=Bk[:, B_INT + 15:].contiguous())


def repartition_tables(t: LocalTables, coord_axis: int, r_ghost: float, world: int, group=None) -> Tuple[LocalTables, torch.Tensor]:
    """One full repartition: bounds -> streams -> exchange -> local tables.  Returns (tables, inner bounds)."""
    own = t.owned
    bounds = slab_bounds(t.state[own, coord_axis].contiguous(), world, group)
    (P, Pd), (B, Bd), (Cs, Cd) = make_streams(t, coord_axis, bounds, r_ghost, world)
    P_in, P_src = all_to_all_rows(P, Pd, world, group)
    B_in, _ = all_to_all_rows(B, Bd, world, group)
    C_in, _ = all_to_all_rows(Cs, Cd, world, group)
    return relocalize(P_in, P_src, B_in, C_in, t.S), bounds


@dataclass
class HaloPlan:
    """Per peer rank: what I send after every step and where what I receive goes (all local indices, int32)."""

    peers: List[int]
    send_idx: Dict[int, torch.Tensor]
    send_rec: Dict[int, torch.Tensor]
    recv_idx: Dict[int, torch.Tensor]
    recv_rec: Dict[int, torch.Tensor]

    def send_doubles(self, r: int) -> int:
        return 9 * len(self.send_idx[r]) + 15 * len(self.send_rec[r])

    def recv_doubles(self, r: int) -> int:
        return 9 * len(self.recv_idx[r]) + 15 * len(self.recv_rec[r])


def build_halo_plan(t: LocalTables, rank: int, world: int, group=None) -> HaloPlan:
    """Request-based: every rank tells the owners which ghosts (by global id) and which of their bond records (by
    (gid_i, gid_j)) it holds; the owners translate the requests into their local indices."""
    dev = t.gid.device
    n = t.n
    ghost = torch.nonzero(~t.owned, as_tuple=True)[0]
    g_src = t.src[ghost]
    # particle requests, grouped by owner rank (stable -> ascending local index inside a group)
    req_p = _i2f(t.gid[ghost])[:, None]
    got_p, got_p_src = all_to_all_rows(req_p, g_src, world, group)
    # bond-record requests: records whose first particle is a ghost
    bptr = t.bptr.to(torch.int64)
    rec_owner = torch.repeat_interleave(torch.arange(n, device=dev), bptr[1:] - bptr[:-1])
    grec = torch.nonzero(~t.owned[rec_owner], as_tuple=True)[0] if t.n_bonds else torch.zeros(0, dtype=torch.int64, device=dev)
    if len(grec):  # only the records an owned particle's bridge reads as its mirror: ghost (j, i) with i owned here
        grec = grec[t.owned[t.bj[grec].to(torch.int64)]]
    r_src = t.src[rec_owner[grec]]
    req_b = torch.stack([_i2f(t.gid[rec_owner[grec]]), _i2f(t.gid[t.bj[grec].to(torch.int64)])], dim=1) if len(grec) else \
        torch.zeros((0, 2), dtype=torch.float64, device=dev)
    got_b, got_b_src = all_to_all_rows(req_b, r_src, world, group)
    # translate what others asked of me
    want_gid = _f2i(got_p[:, 0].contiguous())
    s_idx, ok = lookup_sorted(t.gid, want_gid)
    if len(s_idx) and not bool((ok & t.owned[s_idx]).all()):
        raise RuntimeError("halo plan: a rank requested a particle this rank does not own")
    my_key = rec_owner * max(n, 1) + t.bj.to(torch.int64) if t.n_bonds else torch.zeros(0, dtype=torch.int64, device=dev)
    if len(got_b):
        wi, ok_i = lookup_sorted(t.gid, _f2i(got_b[:, 0].contiguous()))
        wj, ok_j = lookup_sorted(t.gid, _f2i(got_b[:, 1].contiguous()))
        s_rec, ok_r = lookup_sorted(my_key, wi * max(n, 1) + wj)
        if not bool((ok_i & ok_j & ok_r).all()):
            raise RuntimeError("halo plan: a rank requested a bond record this rank does not hold")
    else:
        s_rec = torch.zeros(0, dtype=torch.int64, device=dev)
    peers = sorted(set(g_src.tolist()) | set(got_p_src.tolist()) | set(r_src.tolist()) | set(got_b_src.tolist()))
    peers = [p for p in peers if p != rank]
    order_p = torch.argsort(g_src, stable=True)
    order_b = torch.argsort(r_src, stable=True)
    plan = HaloPlan(peers, {}, {}, {}, {})
    for r in peers:
        plan.send_idx[r] = s_idx[got_p_src == r].to(torch.int32).contiguous()
        plan.send_rec[r] = s_rec[got_b_src == r].to(torch.int32).contiguous()
        gp = ghost[order_p]
        plan.recv_idx[r] = gp[g_src[order_p] == r].to(torch.int32).contiguous()
        gb = grec[order_b]
        plan.recv_rec[r] = gb[r_src[order_b] == r].to(torch.int32).contiguous()
    return plan


def block_tables(rank: int, world: int, x, v, theta, omega, S: int, bond_ij=None, bond_conn=None, bond_len=None,
                 device="cpu", extra_cols: int = 0) -> LocalTables:
    """Initial, arbitrary distribution before the first repartition: rank r 'owns' the contiguous block of global
    ids [r n / W, (r + 1) n / W) of the caller's arrays, without ghosts.  Bond records (global ids, both
    orientations) are taken for the block's particles; partners outside the block are dropped here and come back
    with the first repartition, which ships every record with its particle."""
    n = len(x)
    lo, hi = (rank * n) // world, ((rank + 1) * n) // world
    st = np.concatenate([np.asarray(a, dtype=np.float64)[lo:hi] for a in (x, v, theta, omega)] +
                        [np.zeros((hi - lo, 6 + extra_cols))], axis=1)  # a, alpha, then per-facet wall history (zero)
    gid = torch.arange(lo, hi, dtype=torch.int64, device=device)
    m = hi - lo
    state = torch.from_numpy(st).to(device)
    if bond_ij is not None and len(bond_ij):
        ij = np.asarray(bond_ij, dtype=np.int64)
        sel = (ij[:, 0] >= lo) & (ij[:, 0] < hi)
        B = np.concatenate([ij[sel].view(np.float64), np.asarray(bond_conn)[sel].reshape(-1, 15),
                            np.asarray(bond_len)[sel].reshape(-1, 5)], axis=1)
        B = torch.from_numpy(np.ascontiguousarray(B)).to(device)
    else:
        B = torch.zeros((0, B_INT + B_DBL), dtype=torch.float64, device=device)
    return _InitialTables(gid, state, S, B, m)


class _InitialTables(LocalTables):
    """Block distribution whose bond stream is already in global-id form (partners need not be local)."""

    def __init__(self, gid, state, S, B, m):
        dev = gid.device
        super().__init__(gid=gid, owned=torch.ones(m, dtype=torch.bool, device=dev),
                         src=torch.zeros(m, dtype=torch.int64, device=dev), state=state, S=S,
                         sgid=torch.full((m, S), -1, dtype=torch.int32, device=dev),
                         spr=torch.zeros((m, S, 9), dtype=torch.float64, device=dev),
                         bptr=torch.zeros(m + 1, dtype=torch.int32, device=dev), bj=torch.zeros(0, dtype=torch.int32, device=dev),
                         brev=torch.zeros(0, dtype=torch.int32, device=dev), bconn=torch.zeros((0, 15), dtype=torch.float64, device=dev),
                         blen=torch.zeros((0, 5), dtype=torch.float64, device=dev))
        self.global_bonds = B


# =================================================================================================
# the GPU driver
# =================================================================================================
class _DevView:
    """Zero-copy torch view of device memory owned by the C library (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _view(ptr, shape, dtype: torch.dtype, device) -> torch.Tensor:
    typestr = {torch.float64: "<f8", torch.int32: "<i4", torch.uint8: "|u1"}[dtype]
    if int(np.prod(shape)) == 0 or not ptr:
        return torch.zeros(tuple(shape), dtype=dtype, device=device)
    return torch.as_tensor(_DevView(ptr, shape, typestr), device=device)


class SlabSystem:
    """A granular system slab-decomposed over the ranks of a ``torch.distributed`` process group (one GPU each).

    Mirrors the driver-facing surface of ``GranularSystem`` (do_step, update_neighbor_list, run, get, diagnostics).
    Every rank constructs it with the same full initial state (the driver program's x0, v0, theta0, omega0)."""

    def __init__(self, cfg, x0, v0, theta0, omega0, unit_vectors=None, device: int = 0, group=None, skin: Optional[float] = None,
                 axis: int = 0, max_contacts: int = 16, math_mode: int = 0, bonds=None):
        from . import system as _sys
        from .scenarios import MODEL_CONTACT, MODEL_FACET, MODEL_HAMAKER, MODEL_SINTER

        if not cfg.use_neighbor_list:
            raise ValueError("slab decomposition needs the neighbor-list system (granular_system_neighbor_list)")
        # one triangular facet (surface contact + surface Hamaker) as in DemConfig; its geometry is global, its per-particle
        # wall springs migrate with their particle inside the state rows
        self.n_facets = 1 if cfg.model_mask & MODEL_FACET else 0
        self.facet_vertices = np.ascontiguousarray(np.asarray(cfg.facet_vertices, dtype=np.float64).reshape(3, 3)).copy()
        self._sys = _sys
        self.lib = _sys.load_library()
        self.cfg = cfg
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device("cuda", device)
        self.dev_index = device
        self.axis = axis
        self.skin = float(skin) if skin is not None else 1.0 * cfg.r_part
        self.r_ghost = cfg.r_verlet + self.skin
        self.S = max_contacts
        self.math_mode = int(math_mode)
        self.n_global = len(x0)
        self.h = C.c_void_p()
        self.halo = C.c_void_p()
        self.transport = "none"
        self.steps = 0
        self.repartitions = 0
        self.kicked = False
        self._models = (bool(cfg.model_mask & MODEL_CONTACT), bool(cfg.model_mask & MODEL_SINTER), bool(cfg.model_mask & MODEL_HAMAKER))
        if bonds is None:
            bonds = (None, None, None)
            if cfg.model_mask & MODEL_SINTER:
                if unit_vectors is None:
                    raise ValueError("the sinter model needs unit_vectors")
                if callable(unit_vectors):  # a generator callback can only be consumed by the sequential host scan
                    bonds = _sys.host_bond_table(x0, cfg.r_part, cfg.critical_separation, unit_vectors)
                else:  # device-side bond detection on this rank's GPU (kernels_bonds.cu); same table bit for bit
                    bonds = _sys.device_bond_table(x0, cfg.r_part, cfg.critical_separation, unit_vectors, device=self.dev_index)
        torch.cuda.set_device(self.device)
        self._skin_slots = [(torch.zeros(1, dtype=torch.float64).pin_memory(), torch.cuda.Event()) for _ in range(2)]
        self._skin_prev, self._skin_turn = None, 0
        self.rows_shape = None
        self._stage = None
        t0 = block_tables(self.rank, self.world, x0, v0, theta0, omega0, self.S, *bonds, device=self.device,
                          extra_cols=10 * self.n_facets)
        self._install(t0)

    # -- plumbing ------------------------------------------------------------------------------------
    def _check(self, rc: int):
        if rc != 0:
            raise self._sys.LgError(rc, self.lib.lg_last_error().decode(errors="replace"))

    def _release_halo(self, collective: bool):
        """Teardown order across ranks: everyone unmaps its neighbours' buffers, barrier, then frees its own."""
        if getattr(self, "halo", None) and self.halo.value:
            self.lib.lg_synchronize(self.h)
            self.lib.lg_halo_disconnect(self.halo)
            if collective and self.world > 1 and dist.is_initialized():
                dist.barrier(group=self.group)
            self.lib.lg_halo_destroy(self.halo)
            self.halo = C.c_void_p()

    def close(self, collective: bool = True):
        self._release_halo(collective)
        if getattr(self, "h", None) and self.h.value:
            self.lib.lg_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close(collective=False)
        except Exception:
            pass

    def _install(self, t_old: LocalTables):
        """Repartition `t_old` and build this rank's lg_system from what arrives."""
        t, self.bounds = repartition_tables(t_old, self.axis, self.r_ghost, self.world, self.group)
        torch.cuda.current_stream(self.device).synchronize()  # the library copies from these tensors on its own stream
        self.close()
        lib, cfg = self.lib, self.cfg
        n = t.n
        if n == 0:
            raise RuntimeError("slab decomposition left a rank without particles")
        self._check(lib.lg_create(n, cfg.r_part, cfg.mass, cfg.inertia, self.dev_index, C.byref(self.h)))
        self._check(lib.lg_set_max_contacts(self.h, self.S))
        if self.math_mode:
            self._check(lib.lg_set_math_mode(self.h, self.math_mode))
            if self.rows_shape is not None:
                self._check(lib.lg_set_rows_shape(self.h, self.rows_shape))
        st = t.state
        cols = [st[:, 3 * k:3 * k + 3].contiguous() for k in range(6)]  # x v theta omega a alpha
        owned_u8 = t.owned.to(torch.uint8).contiguous()
        torch.cuda.current_stream(self.device).synchronize()  # the library copies on its own stream: the slices must exist
        self._check(lib.lg_import_state_device(self.h, *[c.data_ptr() for c in cols], int(self.kicked)))
        self._check(lib.lg_set_owned_device(self.h, owned_u8.data_ptr()))
        contact, sinter, hamaker = self._models
        cp = self._sys.ContactParamsStruct.from_seq(cfg.contact)
        if contact:
            self._check(lib.lg_set_contact(self.h, C.byref(cp)))
        if sinter:
            self._check(lib.lg_import_bonds_device(self.h, C.byref(cp), cfg.k_d, cfg.gamma_d, t.n_bonds, t.bptr.data_ptr(),
                                                   t.bj.data_ptr(), t.brev.data_ptr(), t.bconn.data_ptr(), t.blen.data_ptr()))
        if hamaker:
            self._check(lib.lg_set_hamaker(self.h, cfg.hamaker_A, cfg.hamaker_h0))
        self._check(lib.lg_import_slots_device(self.h, t.sgid.contiguous().data_ptr(), t.spr.contiguous().data_ptr()))
        for f in range(self.n_facets):
            # unit normal and area are fixed at construction (triangular_facet.h:26-28) and NOT recomputed when the facet moves:
            # construct from the original vertices, then overwrite the vertices with the current ones
            verts = np.ascontiguousarray(np.asarray(cfg.facet_vertices, dtype=np.float64).reshape(9))
            vf = np.ascontiguousarray(np.asarray(cfg.facet_v, dtype=np.float64).reshape(3))
            sc = self._sys.ContactParamsStruct.from_seq(cfg.surf_contact)
            fid = C.c_int(-1)
            dp = C.POINTER(C.c_double)
            self._check(lib.lg_add_facet(self.h, verts.ctypes.data_as(dp), vf.ctypes.data_as(dp), C.byref(sc), 1, cfg.surf_A, cfg.surf_h0,
                                         C.byref(fid)))
            cur = np.ascontiguousarray(self.facet_vertices.reshape(9))
            self._check(lib.lg_facet_set_vertices(self.h, fid.value, cur.ctypes.data_as(dp)))
            wall = st[:, 18 + 10 * f:18 + 10 * f + 10]
            wspr = wall[:, :9].contiguous()
            wact = (wall[:, 9] != 0.0).to(torch.uint8).contiguous()
            torch.cuda.current_stream(self.device).synchronize()  # the library copies on its own stream: the slices must exist
            self._check(lib.lg_import_wall_device(self.h, fid.value, wspr.data_ptr(), wact.data_ptr()))
        self._check(lib.lg_enable_neighbor_list(self.h, cfg.r_verlet))
        sp = C.c_void_p()
        self._check(lib.lg_get_stream(self.h, C.byref(sp)))
        self.stream = torch.cuda.ExternalStream(sp.value, device=self.device)
        # what stays on the Python side: identity, ownership, halo lists, reference positions for the skin check
        self.n_local = n
        self.gid, self.owned, self.src = t.gid, t.owned, t.src
        self.own_idx = torch.nonzero(t.owned, as_tuple=True)[0]
        self.x_ref = cols[0][self.own_idx].clone()
        self._skin_prev = None  # measurements of an older partition are void
        self.gid_host = None
        self.n_bonds_local = t.n_bonds
        self.plan = build_halo_plan(t, self.rank, self.world, self.group)
        self.send_buf = {r: torch.empty(max(self.plan.send_doubles(r), 1), dtype=torch.float64, device=self.device) for r in self.plan.peers}
        self.recv_buf = {r: torch.empty(max(self.plan.recv_doubles(r), 1), dtype=torch.float64, device=self.device) for r in self.plan.peers}
        self.halo_bytes_per_step = 8 * sum(self.plan.send_doubles(r) for r in self.plan.peers)
        self.repartitions += 1
        self.transport = "nccl" if self.world > 1 else "none"
        if self.world > 1 and self.device.type == "cuda" and os.environ.get("LGB200_HALO", "p2p") == "p2p":
            self._connect_p2p()

    def _connect_p2p(self):
        """Peer-to-peer transport (lgb200.h: lg_halo_*): every rank exports its receive buffer with cudaIpc, the
        neighbours map it, and from then on the pack kernel stores over NVLink straight into the neighbour's memory.
        Falls back to NCCL send/recv (on every rank together) when the mappings cannot be established."""
        lib, plan, peers = self.lib, self.plan, self.plan.peers
        halo = C.c_void_p()
        exports = None
        recv = (C.c_int64 * max(len(peers), 1))(*[plan.recv_doubles(r) for r in peers])
        ok = len(peers) <= 8 and lib.lg_halo_create(self.h, len(peers), recv, C.byref(halo)) == 0
        if ok:
            exports = {}
            for k, r in enumerate(peers):
                ok = ok and lib.lg_halo_set_lists(halo, k, len(plan.send_idx[r]), plan.send_idx[r].data_ptr(), len(plan.send_rec[r]),
                                                  plan.send_rec[r].data_ptr(), len(plan.recv_idx[r]), plan.recv_idx[r].data_ptr(),
                                                  len(plan.recv_rec[r]), plan.recv_rec[r].data_ptr()) == 0
                hb, offs = C.create_string_buffer(64), (C.c_int64 * 4)()
                ok = ok and lib.lg_halo_export(halo, k, hb, offs) == 0
                exports[r] = (hb.raw, list(offs))
        gathered = [None] * self.world
        dist.all_gather_object(gathered, exports if ok else None, group=self.group)
        ok = ok and all(g is not None for g in gathered)
        if ok:
            for k, r in enumerate(peers):
                handle, offs = gathered[r][self.rank]
                ok = ok and lib.lg_halo_connect(halo, k, handle, (C.c_int64 * 4)(*offs)) == 0
        agree = torch.tensor([int(ok)], dtype=torch.int32, device=self.device)
        dist.all_reduce(agree, op=dist.ReduceOp.MIN, group=self.group)
        self.halo = halo
        if int(agree) == 1:
            self.transport = "p2p"
        else:
            if self.rank == 0:
                print("libgran_b200.slab: cudaIpc peer mappings unavailable (%s); halo exchange over NCCL send/recv"
                      % self.lib.lg_last_error().decode(errors="replace"), flush=True)
            if halo.value:
                self._release_halo(collective=True)
            else:
                dist.barrier(group=self.group)
            self.halo = C.c_void_p()

    def _field_view(self, name: str) -> torch.Tensor:
        p = C.c_void_p()
        self._check(self.lib.lg_device_field(self.h, self._sys.FIELDS[name], C.byref(p)))
        return _view(p.value, (self.n_local, 3), torch.float64, self.device)

    def _export(self) -> LocalTables:
        """Snapshot of this rank's tables (copies: the system they live in is about to be replaced)."""
        lib = self.lib
        self._check(lib.lg_flush(self.h))
        S, p_sgid, p_spr = C.c_int(), C.c_void_p(), C.c_void_p()
        self._check(lib.lg_export_slots(self.h, C.byref(S), C.byref(p_sgid), C.byref(p_spr)))
        nb, p_bptr, p_bj, p_bconn, p_blen = C.c_int64(), C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_void_p()
        self._check(lib.lg_export_bonds(self.h, C.byref(nb), C.byref(p_bptr), C.byref(p_bj), C.byref(p_bconn), C.byref(p_blen)))
        self._check(lib.lg_synchronize(self.h))
        n, dev = self.n_local, self.device
        cols = [self._field_view(f) for f in ("x", "v", "theta", "omega", "a", "alpha")]
        for f in range(self.n_facets):  # wall history and the facet's current vertices (moved by update_positions)
            p_ws, p_wa = C.c_void_p(), C.c_void_p()
            self._check(lib.lg_export_wall(self.h, f, C.byref(p_ws), C.byref(p_wa)))
            cols.append(_view(p_ws.value, (n, 9), torch.float64, dev))
            cols.append(_view(p_wa.value, (n,), torch.uint8, dev).to(torch.float64)[:, None])
            verts = np.zeros(9)
            self._check(lib.lg_facet_get_vertices(self.h, f, verts.ctypes.data_as(C.POINTER(C.c_double))))
            self.facet_vertices = verts.reshape(3, 3)
        state = torch.cat(cols, dim=1)
        nbv = int(nb.value)
        return LocalTables(
            gid=self.gid, owned=self.owned, src=self.src, state=state, S=S.value,
            sgid=_view(p_sgid.value, (n, S.value), torch.int32, dev).clone(),
            spr=_view(p_spr.value, (n, S.value, 9), torch.float64, dev).clone(),
            bptr=_view(p_bptr.value, (n + 1,), torch.int32, dev).clone(), bj=_view(p_bj.value, (nbv,), torch.int32, dev).clone(),
            brev=torch.zeros(nbv, dtype=torch.int32, device=dev),
            bconn=_view(p_bconn.value, (nbv, 15), torch.float64, dev).clone(),
            blen=_view(p_blen.value, (nbv, 5), torch.float64, dev).clone())

    # -- halo ----------------------------------------------------------------------------------------------
    def halo_exchange(self):
        """Owners' x, v, omega and bond connectors -> the ghost copies on the neighbouring ranks."""
        plan, lib = self.plan, self.lib
        if not plan.peers:
            return
        if self.transport == "p2p":
            self._check(lib.lg_halo_exchange(self.h, self.halo))
            return
        with torch.cuda.stream(self.stream):
            ops = []
            for r in plan.peers:
                self._check(lib.lg_halo_pack(self.h, len(plan.send_idx[r]), plan.send_idx[r].data_ptr(), len(plan.send_rec[r]),
                                             plan.send_rec[r].data_ptr(), self.send_buf[r].data_ptr()))
                if plan.send_doubles(r):
                    ops.append(dist.P2POp(dist.isend, self.send_buf[r][:plan.send_doubles(r)], r, group=self.group))
                if plan.recv_doubles(r):
                    ops.append(dist.P2POp(dist.irecv, self.recv_buf[r][:plan.recv_doubles(r)], r, group=self.group))
            for q in dist.batch_isend_irecv(ops) if ops else []:
                q.wait()
            for r in plan.peers:
                self._check(lib.lg_halo_unpack(self.h, len(plan.recv_idx[r]), plan.recv_idx[r].data_ptr(), len(plan.recv_rec[r]),
                                               plan.recv_rec[r].data_ptr(), self.recv_buf[r].data_ptr()))

    # -- the driver-facing surface ---------------------------------------------------------------------
    def do_step(self, dt: float):
        if not self.kicked:
            # first call: evaluate, half kick back | refresh the ghosts' kicked v, omega | evaluate again, integrate
            self._check(self.lib.lg_first_half_kick(self.h, float(dt)))
            self.halo_exchange()
            self.kicked = True
        self._check(self.lib.lg_do_step(self.h, float(dt)))
        self.halo_exchange()
        self.steps += 1

    def needs_repartition(self) -> bool:
        """True on every rank when some owned particle has moved too far since the last partition.

        The measurement (max displacement, all-reduced) is queued on the system's stream and copied to pinned memory
        without waiting; what is consumed here is the measurement of the PREVIOUS rebuild, which has long arrived -- no host
        synchronisation on the every-20-steps path.  Being one period late is paid for with a halved threshold (skin / 4
        instead of skin / 2): the ghost margin is violated only if the displacement more than doubles within one period.
        The first check after a (re)partition has nothing to consume and waits."""
        with torch.cuda.stream(self.stream):
            x = self._field_view("x")[self.own_idx]
            d2 = ((x - self.x_ref) ** 2).sum(dim=1).max().reshape(1)
            if self.world > 1:
                dist.all_reduce(d2, op=dist.ReduceOp.MAX, group=self.group)
            ready = torch.cuda.Event()
            ready.record(self.stream)
        # the copy into pinned memory runs on torch's current stream (pinned blocks remember the streams that used them, and
        # the library's stream dies with the system -- see set_state_host), ordered after the measurement by an event
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(ready)
        slot = self._skin_slots[self._skin_turn & 1]
        slot[0].copy_(d2, non_blocking=True)
        slot[1].record(cur)
        prev = self._skin_prev
        self._skin_prev = slot
        self._skin_turn += 1
        if prev is None:  # nothing measured yet for this partition
            slot[1].synchronize()
            return float(slot[0]) > (0.5 * self.skin) ** 2
        prev[1].synchronize()  # recorded a whole period ago
        return float(prev[0]) > (0.25 * self.skin) ** 2

    def repartition(self):
        with torch.cuda.stream(self.stream):
            t = self._export()
        self.stream.synchronize()
        self._install(t)  # on torch's current stream: the system (and its stream) is replaced inside

    def update_neighbor_list(self):
        if self.world > 1 and self.needs_repartition():
            self.repartition()
        self._check(self.lib.lg_update_neighbor_list(self.h))
        if self.math_mode and self.rows_shape is None:
            # tile shape of step_rows2_kernel from GLOBAL statistics (mean row length over all ranks, total particle count),
            # pinned on every rank for the life of the run: every row is then summed in the same order as on one GPU
            out = (C.c_int64 * 8)()
            self._check(self.lib.lg_get_counters(self.h, out))
            pairs = torch.tensor([out[0]], dtype=torch.int64, device=self.device)
            if self.world > 1:
                dist.all_reduce(pairs, group=self.group)
            self.rows_shape = int(self.lib.lg_rows_shape_rule(float(int(pairs)) / max(self.n_global, 1), self.n_global, int(self._models[1])))
            if self.rows_shape == 0 and self.n_global // max(self.world, 1) < 300_000:
                # same lanes per particle (= same summation order as one GPU), half-size tiles: with ~100 k particles per rank
                # 96-particle tiles give each resident CTA 3.2 tiles, i.e. a last wave that is a quarter full
                self.rows_shape = 3
            self._check(self.lib.lg_set_rows_shape(self.h, self.rows_shape))

    def facet_update_positions(self, dt: float, facet_id: int = 0):
        self._check(self.lib.lg_facet_update_positions(self.h, int(facet_id), float(dt)))

    def run(self, dt: float, n_steps: int, rebuild_every: int = 0, move_facets: bool = False):
        k = 0
        while k < n_steps:
            if rebuild_every > 0 and self.steps % rebuild_every == 0:
                self.update_neighbor_list()
            if self.transport == "p2p" and self.kicked:
                # the whole stretch up to the next rebuild is queued by ONE library call (step kernel, push, pull per step)
                chunk = n_steps - k if rebuild_every <= 0 else min(n_steps - k, rebuild_every - self.steps % rebuild_every)
                self._check(self.lib.lg_run_slab(self.h, self.halo, float(dt), chunk, int(move_facets)))
                self.steps += chunk
                k += chunk
            else:
                self.do_step(dt)
                if move_facets:
                    for f in range(self.n_facets):
                        self.facet_update_positions(dt, f)
                k += 1

    def synchronize(self):
        self._check(self.lib.lg_synchronize(self.h))

    def get(self, name: str) -> np.ndarray:
        """Field in the caller's original particle order, assembled on every rank."""
        if name in ("a", "alpha"):
            self._check(self.lib.lg_flush(self.h))
        with torch.cuda.stream(self.stream):
            mine = torch.cat([_i2f(self.gid[self.own_idx])[:, None], self._field_view(name)[self.own_idx]], dim=1)
            if self.world == 1:
                rows = mine
            else:
                dest = torch.arange(self.world, device=self.device).repeat_interleave(len(mine))
                rows, _ = all_to_all_rows(mine.repeat(self.world, 1), dest, self.world, self.group)
            out = torch.empty((self.n_global, 3), dtype=torch.float64, device=self.device)
            out[_f2i(rows[:, 0].contiguous())] = rows[:, 1:]
        self.synchronize()
        torch.cuda.synchronize(self.device)
        return out.cpu().numpy()

    def set_state_host(self, x, v, theta, omega):
        """lg_set_state for a decomposed system.  Every rank copies the id range it holds (owned and ghost rows) of the
        caller's arrays host -> device and picks its rows there; re-arms the first-step half kick like the reference
        constructor.  The copies run on torch's current stream, not
        on the library's: pinned host blocks remember the streams that used them, and the library's stream dies with
        the system."""
        cur = torch.cuda.current_stream(self.device)
        if self.gid_host is None:  # cached per partition: the id range this rank holds
            self.gid_host = (int(self.gid.min()), int(self.gid.max()))
        gmin, gmax = self.gid_host
        rel = self.gid - gmin
        loc = []
        for a in (x, v, theta, omega):
            # only the id range this rank holds (owned + ghosts) crosses PCIe -- one contiguous slice of the caller's array
            # (asynchronous when it is pinned), its rows picked on the device.  Slabs of an input whose ids run along the slab
            # axis span about 1 / world of the array; an input without such order degrades to the full copy.
            t = a if isinstance(a, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(np.asarray(a, dtype=np.float64)))
            loc.append(t.reshape(-1, 3)[gmin:gmax + 1].to(self.device, non_blocking=True).index_select(0, rel).contiguous())
        self.last_upload_rows = gmax - gmin + 1
        x_ref = loc[0][self.own_idx].clone()
        cur.synchronize()
        self._check(self.lib.lg_import_state_device(self.h, *[c.data_ptr() for c in loc], None, None, 0))  # synchronises
        self.kicked = False
        self.x_ref = x_ref

    def get_owned_host(self, names) -> dict:
        """D2H of this rank's owned rows (global ids + fields) into pinned buffers, without assembling the global arrays."""
        out = {"gid": self.gid[self.own_idx].cpu().numpy()}
        src = {}
        with torch.cuda.stream(self.stream):
            for f in names:
                if f in ("a", "alpha"):
                    self._check(self.lib.lg_flush(self.h))
                src[f] = self._field_view(f)[self.own_idx]
        self.stream.synchronize()
        for f in names:  # on torch's current stream (see set_state_host)
            buf = torch.empty(src[f].shape, dtype=torch.float64, pin_memory=True)
            buf.copy_(src[f], non_blocking=True)
            out[f] = buf.numpy()
        torch.cuda.current_stream(self.device).synchronize()
        return out

    def counters(self) -> dict:
        out = (C.c_int64 * 8)()
        self._check(self.lib.lg_get_counters(self.h, out))
        loc = torch.tensor([out[0], out[1], int(self.owned.sum()), self.n_local - int(self.owned.sum())], dtype=torch.int64,
                           device=self.device)
        tot = loc.clone()
        if self.world > 1:
            dist.all_reduce(tot, group=self.group)
        return {"pairs": int(tot[0]), "contacts": int(tot[1]), "owned_local": int(loc[2]), "ghosts_local": int(loc[3]),
                "ghosts_total": int(tot[3]), "launches": int(out[5]), "repartitions": self.repartitions,
                "halo_bytes_per_step": self.halo_bytes_per_step, "transport": self.transport}

    def _partial(self, pass_: int, com: Optional[torch.Tensor] = None) -> torch.Tensor:
        """lg_diagnostics_partial: the device reduction of lg_diagnostics over this rank's OWNED particles, left on the
        device (call inside the system's stream context)."""
        p = C.c_void_p()
        self._check(self.lib.lg_diagnostics_partial(self.h, int(pass_), com.data_ptr() if com is not None else None, C.byref(p)))
        return _view(p.value, (12,), torch.float64, self.device).clone()

    def kinetic_energy(self) -> float:
        """compute_total_kinetic_energy (test/compute_energy.cpp:7-17): lg_diagnostics partial sums of the owned particles,
        all-reduced over the ranks."""
        with torch.cuda.stream(self.stream):
            s = self._partial(0)[:2]
            if self.world > 1:
                dist.all_reduce(s, group=self.group)
            return float(s.sum())  # inside the stream context, see needs_repartition

    def diagnostics(self) -> dict:
        """test/compute_energy.cpp:7-58, test/mass_distribution.cpp:7-26: lg_diagnostics partial sums -> all_reduce(SUM)."""
        cfg = self.cfg
        with torch.cuda.stream(self.stream):
            s = self._partial(0)[:11]
            if self.world > 1:
                dist.all_reduce(s, group=self.group)
            com = (s[8:11] / self.n_global).contiguous()
            r = self._partial(1, com)[11:12]
            if self.world > 1:
                dist.all_reduce(r, group=self.group)
            s, rg2, com = s.cpu().numpy(), float(r), com.cpu().numpy()
        return {"ke": s[0] + s[1], "ke_trs": s[0], "ke_rot": s[1], "linear_momentum": cfg.mass * float(np.linalg.norm(s[2:5])),
                "angular_momentum": float(np.linalg.norm(s[5:8])), "com": com, "rg": float(np.sqrt(rg2 / self.n_global))}
