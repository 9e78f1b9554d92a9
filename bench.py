#!/usr/bin/env python
"""bench.py -- particle-timesteps/s of the DEM hot path on BASELINE.json's configs (default: config 4, the headline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config 2|3|4|4c|5|5d]

  --config 4   (default, configs[3], the one `metric` is quoted on) 1 000 050-particle multi-aggregate collision
  --config 4c  config 4 at the moment of impact: rows of aggregates slid into first contact (Luding contacts > 0)
  --config 2   (configs[1]) 10 000-particle fractal aggregate, contact + Hamaker; the reference arm runs the
               UNMODIFIED reference at full size (7.2 GB dense spring table) -- the one same-config comparison
  --config 3   (configs[2]) 100 000-particle sintered fractal aggregate on a moving facet (CPU leg: oracle port)
  --config 5   (configs[4]) random packing, volume fraction 0.30, contact + Hamaker, 1.25 M particles PER GPU (weak
               scaling sweep over --gpus)
  --config 5d  dense jittered lattice (rows of ~80 neighbors, ~4 contacts per particle), 1.25 M particles per GPU:
               the pair-tile kernel's regime

Workload (config.workload): ~1M-particle multi-aggregate collision -- 6667 randomly rotated copies of the
150-particle soot aggregate (resources/aggregate.h), left half moving +x, right half -x; sinter bridges with
embedded Luding contact + Hamaker, neighbor list r_verlet = 5 r rebuilt every 20 steps (SURVEY.md 8d cfg 4).
One bench "step" = one DEM time step (do_step) over all particles; the neighbor rebuild that falls on every
20th step is inside the timed region.  Prints ONE JSON line (see the task contract).

  value     whole-job particle-timesteps/s with state resident in HBM (lg_run, CUDA events on the system stream)
  e2e       the same through the host-facing API with HOST buffers: lg_set_state from pinned memory (H2D),
            K x [lg_do_step + lg_diagnostics read-back], final get_x/get_v/get_omega/get_theta (D2H)
  roofline  the pair-force kernel: algorithmic bytes per launch / mean CUDA-event duration of its launches
  cpu_baseline / --impl reference
            the reference's own OpenMP implementation (oracle/_ref/libref_timing.so, built from the unmodified
            reference with its -O3 flags) on a bounded sample of the same workload (the first `sample` aggregates;
            the full size needs 232 B * N^2 = 232 TB of dense tables and cannot run), all host threads.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
from ctypes import byref, c_double as C_double, c_int64 as C_int64, create_string_buffer

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from libgran_b200 import scenarios as S  # noqa: E402

METRIC = "particle-timesteps/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, read at bench time from the summary of
# the most recent `ncu --set full` capture of the same command (scripts/ncu_traffic.py writes profiles/traffic_ncu.json:
# {"<config>:<math>": {"kernel": ..., "bytes": ..., "source": "<raw csv>"}}); None when no capture of this workload exists
TRAFFIC_FILE = os.path.join(ROOT, "profiles", "traffic_ncu.json")
REBUILD_EVERY = int(os.environ.get("LGB200_BENCH_REBUILD_EVERY", "20"))  # 20 = the reference drivers' period; the override is for diagnostics only


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples SM clocks and throttle reasons of one GPU during the timed region: NVML polled every 2 ms from a
    thread (the timed region lasts tens of milliseconds, far less than nvidia-smi's shortest loop period)."""

    def __init__(self, gpu_index: int = 0):
        self.gpu = gpu_index
        self.sm, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self.thread = None
        self.nv = None

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            self.nv = nv
            self.handle = nv.nvmlDeviceGetHandleByIndex(self.gpu)
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception:
            self.nv = None

    def _poll(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        while True:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                for name, bit in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            if self._stop.wait(0.002):
                break

    def stop(self):
        if self.nv is None or self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        self._stop.set()
        self.thread.join(timeout=2)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.sm)}


def numpy_unit_vectors(count: int, seed: int = 99) -> np.ndarray:
    rng = np.random.Generator(np.random.MT19937(seed))
    v = rng.normal(size=(count, 3))
    return v / np.linalg.norm(v, axis=1)[:, None]


def ncu_traffic(config: str, math: str, kernel: str):
    try:
        with open(TRAFFIC_FILE) as f:
            rec = json.load(f).get(f"{config}:{math}")
        if rec and kernel.startswith(rec["kernel"]):
            return float(rec["bytes"])
    except Exception:
        pass
    return None


# ---- workloads -----------------------------------------------------------------------------------------
PER_GPU_5 = 1_250_000


def build_workload(args, world: int = 1, sample: bool = False):
    """(scenario, description, move_facets) of --config; `sample` = the bounded CPU-reference sample of the same workload."""
    c = args.config
    if c == "2":
        n = 10_000
        return S.fractal_aggregate(n), f"cfg2 fractal aggregate, {n} particles, contact+Hamaker", False
    if c == "3":
        n = 100_000
        return (S.sintered_aggregate_on_facet(n),
                f"cfg3 sintered fractal aggregate on a facet moving +z at 0.5 m/s, {n} particles, sinter+contact+Hamaker+wall contact+wall Hamaker", True)
    if c in ("4", "4c"):
        copies = args.sample if sample else args.copies
        if c == "4":
            scn = S.multi_aggregate_collision(copies)
            return scn, (f"cfg4 multi-aggregate collision, {scn.cfg.n} particles ({copies} x aggregate.h), sinter+contact+Hamaker "
                         "(the halves approach at 1 m/s: no aggregate touches another inside the timed window, contacts = 0 -- bridges + "
                         "Hamaker work; --config 4c is the same system at impact)"), False
        # CPU sample: the separated layout -- the reference's sinter constructor would bond the touching aggregates together
        scn = S.multi_aggregate_collision(copies) if sample else S.multi_aggregate_in_contact(copies)
        return scn, (f"cfg4c multi-aggregate collision at impact (rows of aggregates slid into first contact at the static-equilibrium overlap), "
                     f"{scn.cfg.n} particles ({copies} x aggregate.h), sinter+contact+Hamaker"), False
    if c == "5":
        n = 125_000 if sample else PER_GPU_5 * world
        scn = S.random_packing(n, volume_fraction=0.30)
        return scn, f"cfg5 random packing (RSA, volume fraction 0.30), {n} particles, contact+Hamaker", False
    if c == "5d":
        n1 = 50 if sample else 108
        nx = n1 * (1 if sample else world)
        scn = S.lattice_block(nx, n1, n1, spacing=1.98, jitter=0.02, v_scale=0.05)
        return scn, f"cfg5d dense jittered lattice {nx}x{n1}x{n1} (spacing 1.98 r), {scn.cfg.n} particles, contact+Hamaker", False
    raise SystemExit(f"unknown --config {c}")


def morton_presort(scn):
    """The same particles listed along a Morton curve (cells of one diameter).  The slab driver keeps the caller's global-id
    order on every rank (that is what makes a decomposed run bit-identical to one GPU), so a synthetic input whose ids carry
    no locality (random packing, fractal cluster) is handed to it spatially ordered; the single-GPU arm leaves the caller's
    order alone and renumbers internally (lg_set_reorder)."""
    x = scn.x0
    c = np.floor((x - x.min(axis=0)) / (2.0 * scn.cfg.r_part)).astype(np.uint64)
    c = np.minimum(c, 1023)

    def spread(v):
        v = (v | (v << np.uint64(16))) & np.uint64(0x030000FF)
        v = (v | (v << np.uint64(8))) & np.uint64(0x0300F00F)
        v = (v | (v << np.uint64(4))) & np.uint64(0x030C30C3)
        v = (v | (v << np.uint64(2))) & np.uint64(0x09249249)
        return v

    key = np.zeros(len(x), dtype=np.uint64)
    for axis in range(3):  # 10 bits per axis: two 5-bit halves through the 8-bit spreader
        lo, hi = c[:, axis] & np.uint64(31), c[:, axis] >> np.uint64(5)
        key |= (spread(lo) << np.uint64(axis)) | (spread(hi) << np.uint64(15 + axis))
    order = np.argsort(key, kind="stable")
    return S.Scenario(scn.name + "_morton", scn.cfg, *(np.ascontiguousarray(a[order]) for a in scn.state))


def workload_line(desc: str) -> str:
    return f"{desc}, r_verlet=5r, rebuild every {REBUILD_EVERY} steps"


def effective_warmup(args) -> int:
    """Untimed steps before the timed region: at least the requested W, at least 3, and enough to pass two neighbor
    rebuilds (the second one is the first that takes the steady-state path and sizes its buffers)."""
    return max(args.warmup, 3, 2 * REBUILD_EVERY + 1)


def algorithmic_bytes(n, pairs, contacts, bonds_ordered):
    """bytes per DEM step (cache-perfect, FP64, int32 indices) and per force-kernel launch."""
    step = 192 * n + 4 * (n + 1) + 4 * pairs + 144 * contacts + 280 * bonds_ordered
    # the pair-force kernel alone: reads x, v, omega (72 N), writes a, alpha (48 N), row pointers, columns,
    # contact springs r+w, own connectors (120) + rest lengths (40) per ordered bond
    force = 72 * n + 48 * n + 4 * (n + 1) + 4 * pairs + 144 * contacts + 160 * bonds_ordered
    return step, force


def run_ours(args):
    import torch
    import torch.distributed as dist

    from libgran_b200.system import from_config

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    scn, desc, moves = build_workload(args, world)
    cfg = scn.cfg
    n = cfg.n
    if world > 1:
        if args.config in ("2", "3", "5", "5d") and args.reorder != "off":
            scn = morton_presort(scn)
            desc += ", input listed along a Morton curve (host pre-sort; the slab driver keeps the caller's id order)"
        return run_slab(args, scn, desc, moves, world, rank, local_rank)

    uv = numpy_unit_vectors((4 * n if args.config == "3" else n) + 1024) if cfg.model_mask & S.MODEL_SINTER else None
    mode = 1 if (args.math == "fast" and args.variant in (None, 1, 2)) else 0
    bonds = scn.make_bonds(uv) if scn.make_bonds is not None else None
    reorder = (args.reorder == "on") or (args.reorder == "auto" and mode == 1 and args.config in ("2", "3", "5", "5d"))
    sysm = from_config(cfg, *scn.state, unit_vectors=uv, device=local_rank, variant=args.variant, math_mode=mode, bonds=bonds, reorder=reorder)
    if mode == 1 and not args.no_autotune:
        sysm.set_autotune(True)  # tile shape of step_rows2_kernel measured on the first fused launches (inside the warm-up)
    stream = torch.cuda.ExternalStream(sysm.stream_handle(), device=torch.device("cuda", local_rank))

    def timed_run(k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            e0.record()
            sysm.run(cfg.dt, k, REBUILD_EVERY, moves)
            e1.record()
        sysm.synchronize()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    # ---- device-resident throughput ------------------------------------------------------------
    sysm.run(cfg.dt, effective_warmup(args), REBUILD_EVERY, moves)
    sysm.synchronize()
    c0 = sysm.counters()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = timed_run(args.steps)
    clocks = sampler.stop()
    c1 = sysm.counters()
    value = n * args.steps / (ms * 1e-3)
    launches = c1["launches"] - c0["launches"]

    # ---- roofline of the dominant kernel (the step kernel), measured live with events ----------------
    sysm.set_profiling(True)
    prof_steps = min(args.steps, 40)
    sysm.run(cfg.dt, prof_steps, REBUILD_EVERY, moves)
    force_ms, force_launches = sysm.get_profile()
    sysm.set_profiling(False)
    kernel = sysm.last_step_kernel()
    cnt = sysm.counters()
    step_bytes, force_bytes = algorithmic_bytes(n, cnt["pairs"], cnt["contacts"], cnt["bonds"])
    peak, peak_src = measured_peaks()
    force_avg_ms = force_ms / max(force_launches, 1)
    # the production launch is the FUSED step kernel (force + Verlet + connector update): it moves the whole B_step;
    # the validation variant's force-only kernel moves force_bytes
    fused = args.variant in (None, 1, 2)
    launch_bytes = step_bytes if fused else force_bytes
    achieved = launch_bytes / (force_avg_ms * 1e-3) / 1e9
    kname = kernel + (" (fused pair force + Verlet + connectors)" if fused else " (pair force)")
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": ncu_traffic(args.config, args.math, kernel), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(launch_bytes), "avg_launch_ms": round(force_avg_ms, 5),
                "launches_timed": int(force_launches),
                "share_of_step": round(force_avg_ms / (ms / args.steps), 3),
                "whole_step": {"algorithmic_bytes_per_step": int(step_bytes),
                               "achieved_gbs": round(step_bytes * args.steps / (ms * 1e-3) / 1e9, 1),
                               "frac": round(step_bytes * args.steps / (ms * 1e-3) / 1e9 / peak, 4),
                               "frac_of_nominal_8TBs": round(step_bytes * args.steps / (ms * 1e-3) / 8e12, 4)}}

    if args.kernel_only:
        if rank == 0:
            print(json.dumps({"config": args.config, "math": args.math, "value": round(value, 1), "ms_per_step": round(ms / args.steps, 5),
                              "particles": n, "pairs": cnt["pairs"], "contacts": cnt["contacts"], "bonds_ordered": cnt["bonds"],
                              "roofline": roofline, "gpu_launches": int(launches), "clocks": clocks,
                              "reorder": reorder, "rows_shape": sysm.rows_shape(), "lib": os.environ.get("LGB200_LIB", "default"), "tag": os.environ.get("LGB200_TAG")}))
        return

    # ---- end to end through the host-facing API with host buffers ---------------------------------
    pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in scn.state]
    host_np = [t.numpy() for t in pinned]
    outs = [torch.empty((n, 3), dtype=torch.float64).pin_memory() for _ in range(4)]
    out_np = [t.numpy() for t in outs]
    e2e_sys = from_config(cfg, *scn.state, unit_vectors=uv, device=local_rank, variant=args.variant, math_mode=mode, bonds=bonds, reorder=reorder)  # model tables built once (as a ctor would)
    if mode == 1 and not args.no_autotune:
        e2e_sys.set_autotune(True)
    facet0 = e2e_sys.facet_vertices(0) if moves else None
    e2e_steps = args.steps
    # untimed priming with dt = 0: sizes the neighbor / scratch buffers for exactly this state and leaves positions,
    # velocities and bridge connectors where the constructor put them (x += v * 0, c += 0)
    e2e_sys.run(0.0, effective_warmup(args), REBUILD_EVERY)
    e2e_sys.synchronize()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_sys.set_state(*host_np)  # H2D of x0, v0, theta0, omega0 (a new trajectory: history springs and rows reset)
    ke = 0.0
    for k in range(e2e_steps):
        if k % REBUILD_EVERY == 0:
            e2e_sys.update_neighbor_list()
        e2e_sys.do_step(cfg.dt)
        if moves:
            e2e_sys.facet_update_positions(cfg.dt)
        if (k + 1) % 10 == 0:
            ke = e2e_sys.diagnostics()["ke"]  # D2H of the step's metric (as main.cpp:152-159 dumps energies)
    for name, buf in zip(("x", "v", "omega", "theta"), out_np):
        e2e_sys.get(name, out=buf)  # D2H
    e2e_sys.synchronize()
    t1 = time.perf_counter()
    e2e_value = n * e2e_steps / (t1 - t0)
    h2d = 4 * 24 * n / e2e_steps
    d2h = (4 * 24 * n + (e2e_steps // 10) * 9 * 8) / e2e_steps
    e2e = {"value": round(e2e_value, 1), "unit": METRIC, "h2d_bytes_per_step": round(h2d, 1),
           "d2h_bytes_per_step": round(d2h, 1), "steps": e2e_steps, "final_ke": ke,
           "what": "lg_set_state(H2D, pinned) + K x lg_do_step (+ rebuild/20, KE read-back/10) + get x,v,omega,theta (D2H)"}
    del facet0
    e2e_sys.close()

    line = {
        "metric": METRIC, "value": round(value, 1), "unit": METRIC, "n_gpus": world, "steps": args.steps,
        "warmup": effective_warmup(args), "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True,
        "scaling": "weak" if args.config in ("5", "5d") else "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_line(desc), "config": args.config, "particles": n, "pairs": cnt["pairs"],
                   "bonds_ordered": cnt["bonds"], "contacts": cnt["contacts"],
                   "l2_policy": ("working set larger than L2" if step_bytes > 3 * 126e6 else
                                 f"working set {step_bytes / 1e6:.1f} MB per step is L2-resident by nature of this config (the whole "
                                 "system is smaller than the 126 MB L2); every byte is rewritten each step"),
                   "kernel_variant": args.variant, "internal_renumbering": reorder, "rows_tile_shape": sysm.rows_shape(), "math_mode": "fast (<=1e-10 rel. forces, bit-exact sets)" if mode else "exact (bit-identical to the reference)"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
    }
    if rank == 0 and not args.no_cpu_baseline:
        line["cpu_baseline"] = run_reference_sample(args, budget_s=args.cpu_seconds, in_gpu_line=True)
    if rank == 0:
        print(json.dumps(line))


def run_slab(args, scn, desc, moves, world, rank, local_rank):
    """N > 1: the same workload slab-decomposed along x, one rank per GPU (libgran_b200/slab.py).  Configs 2-4: strong
    scaling (the particle count is fixed); configs 5 / 5d: weak scaling (1.25 M particles per GPU).
    `value` = all particles * steps / max-over-ranks device time."""
    import torch
    import torch.distributed as dist

    from libgran_b200.slab import SlabSystem

    cfg = scn.cfg
    n = cfg.n
    dev = torch.device("cuda", local_rank)
    uv = numpy_unit_vectors((4 * n if args.config == "3" else n) + 1024) if cfg.model_mask & S.MODEL_SINTER else None
    mode = 1 if args.math == "fast" else 0
    bonds = scn.make_bonds(uv) if scn.make_bonds is not None else None
    sysm = SlabSystem(cfg, *scn.state, unit_vectors=uv, device=local_rank, math_mode=mode, bonds=bonds)
    warm = effective_warmup(args)
    sysm.run(cfg.dt, warm, REBUILD_EVERY)
    sysm.synchronize()
    c0 = sysm.counters()

    def timed(k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()
        with torch.cuda.stream(sysm.stream):
            e0.record()
        sysm.run(cfg.dt, k, REBUILD_EVERY)
        with torch.cuda.stream(sysm.stream):
            e1.record()
        sysm.synchronize()
        torch.cuda.synchronize()
        dist.barrier()
        t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        torch.cuda.synchronize()
        return float(t)

    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = timed(args.steps)
    clocks = sampler.stop()
    c1 = sysm.counters()
    value = n * args.steps / (ms * 1e-3)
    launches = torch.tensor([c1["launches"] - c0["launches"]], dtype=torch.int64, device=dev)
    dist.all_reduce(launches)

    # roofline of this rank's step kernel (events around each launch), reported for the slowest rank
    sysm.lib.lg_set_profiling(sysm.h, 1)
    prof_steps = min(args.steps, 40)
    sysm.run(cfg.dt, prof_steps, REBUILD_EVERY)
    fm, fl = C_double(), C_int64()
    sysm.lib.lg_get_profile(sysm.h, byref(fm), byref(fl))
    sysm.lib.lg_set_profiling(sysm.h, 0)
    cnt = sysm.counters()
    own = cnt["owned_local"]
    loc = (C_int64 * 8)()
    sysm.lib.lg_get_counters(sysm.h, loc)
    step_bytes_local = algorithmic_bytes(own, int(loc[0]), int(loc[1]), sysm.n_bonds_local)[0]
    force_avg_ms = fm.value / max(fl.value, 1)
    stat = torch.tensor([force_avg_ms, step_bytes_local / (force_avg_ms * 1e-3) / 1e9], dtype=torch.float64, device=dev)
    worst = stat.clone()
    dist.all_reduce(worst, op=dist.ReduceOp.MAX)
    slow = float(worst[0])
    peak, peak_src = measured_peaks()
    ach = step_bytes_local / (force_avg_ms * 1e-3) / 1e9
    kbuf = create_string_buffer(64)
    sysm.lib.lg_last_step_kernel(sysm.h, kbuf, 64)
    roofline = {"bound": "hbm", "kernel": kbuf.value.decode() + " (fused pair force + Verlet + connectors), this rank", "achieved": round(ach, 1),
                "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(step_bytes_local), "avg_launch_ms": round(force_avg_ms, 5),
                "slowest_rank_launch_ms": round(slow, 5), "launches_timed": int(fl.value),
                "share_of_step": round(force_avg_ms / (ms / args.steps), 3)}

    # end to end: host buffers in, host buffers out, through the slab driver
    # (a fresh system, primed with dt = 0 like the single-GPU arm: same buffers, untouched state and connectors)
    sysm.close()
    sysm = SlabSystem(cfg, *scn.state, unit_vectors=uv, device=local_rank, math_mode=mode, bonds=bonds)
    sysm.run(0.0, warm, REBUILD_EVERY)
    own = sysm.counters()["owned_local"]
    pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in scn.state]
    sysm.get_owned_host(("x", "v", "omega", "theta"))  # untimed: allocates the pinned download buffers once
    e2e_steps = args.steps
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    sysm.set_state_host(*pinned)  # H2D of the caller's arrays from pinned memory, rows of this rank gathered on the device
    ke = 0.0
    sysm.steps = 0  # new positions: the first run() call below rebuilds the neighbor rows
    done = 0
    while done < e2e_steps:
        chunk = min(10, e2e_steps - done)
        sysm.run(cfg.dt, chunk, REBUILD_EVERY)  # the driver-facing loop call (rebuild every 20 steps inside)
        done += chunk
        ke = sysm.kinetic_energy()  # all-reduce + D2H of the step's metric
    outs = sysm.get_owned_host(("x", "v", "omega", "theta"))  # D2H of the owned particles
    sysm.synchronize()
    torch.cuda.synchronize()
    dist.barrier()
    t1 = time.perf_counter()
    tt = torch.tensor([t1 - t0], dtype=torch.float64, device=dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    nloc = torch.tensor([getattr(sysm, "last_upload_rows", n), own], dtype=torch.int64, device=dev)
    dist.all_reduce(nloc)
    e2e = {"value": round(n * e2e_steps / float(tt), 1), "unit": METRIC,
           "h2d_bytes_per_step": round(4 * 24 * int(nloc[0]) / e2e_steps, 1),  # every rank uploads the id range it holds (owned + ghosts)
           "d2h_bytes_per_step": round((4 * 24 * int(nloc[1]) + (e2e_steps // 10) * 12 * 8 * world) / e2e_steps, 1),
           "steps": e2e_steps, "final_ke": ke,
           "what": "per rank: set_state_host (H2D, pinned) + K/10 x [run(10 steps: step + halo exchange, rebuild/20) + KE "
                   "all-reduce + read-back] + owned x,v,omega,theta (D2H)"}
    del outs
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True,
        "scaling": "weak" if args.config in ("5", "5d") else "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_line(desc), "config": args.config, "particles": n, "pairs": cnt["pairs"],
                   "contacts": cnt["contacts"], "math_mode": "fast" if mode else "exact", "l2_policy": "per-rank working set larger than L2 up to 4 GPUs; at 8 GPUs "
                   "(~100 MB/step/rank) the step streams through L2 once per step, no reuse across steps is possible "
                   "because every byte is rewritten each step",
                   "decomposition": f"{world} slabs along x, ghosts {cnt['ghosts_total']}, halo {cnt['halo_bytes_per_step']} B/step "
                                    f"sent by rank 0, repartitions {cnt['repartitions']}, halo transport {cnt['transport']}"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
    }
    if rank == 0:
        print(json.dumps(line))
    sysm.close()
    dist.destroy_process_group()


def host_threads() -> int:
    """Physical cores of the host (the reference's advice: one OpenMP thread per core, Writerside/topics/Installation.md:22-24)."""
    try:
        import psutil

        c = psutil.cpu_count(logical=False)
        if c:
            return int(min(c, len(os.sched_getaffinity(0))))
    except Exception:
        pass
    return max(1, len(os.sched_getaffinity(0)))


def run_reference_sample(args, budget_s: float, in_gpu_line: bool = False):
    """The reference's own CPU implementation of the path on the host cores.  kind "reference" = the unmodified reference
    (oracle/_ref/libref_timing.so, built from the reference's headers with its -O3 -fopenmp flags); kind "port" = the
    oracle restatement, where the reference's dense N x N tables cannot hold the config.  Timed: at least 200 steps and
    at least 2 s (whatever --steps says), after a 20-step warm-up, constructor excluded (test/hamaker_neighbor.cpp:119-127
    times the loop only); the 1-thread figure is taken on a quarter of the steps."""
    os.environ.setdefault("OMP_PROC_BIND", "close")
    os.environ.setdefault("OMP_WAIT_POLICY", "passive")
    from oracle import bindings as ob

    scn, desc, moves = build_workload(args, 1, sample=True)
    cfg = scn.cfg
    use_ref = args.config in ("2", "4", "4c") and ob.ref_available(timing=True)
    threads = host_threads()
    if use_ref:
        kind, lib = "reference", ob.load_ref(timing=True)
        lib.ref_set_threads(threads)   # explicit: under torchrun OMP_NUM_THREADS defaults to 1
        make = lambda: ob.RefSystem(cfg, *scn.state, timing=True)
    else:
        kind, lib = "port", ob.load_oracle()
        lib.orc_set_threads(threads)
        make = lambda: ob.OracleSystem(cfg, *scn.state)
    set_threads = lib.ref_set_threads if use_ref else lib.orc_set_threads

    def loop(sysm, k):
        if not moves:
            sysm.run(cfg.dt, k, REBUILD_EVERY)
            return
        for s in range(k):  # driver loop with facet.update_positions(dt), test/surface.cpp style
            if s % REBUILD_EVERY == 0:
                sysm.update_neighbor_list()
            sysm.do_step(cfg.dt)
            sysm.facet_update_positions(cfg.dt)

    sysm = make()
    loop(sysm, 20)
    t0 = time.perf_counter()
    loop(sysm, 20)
    per20 = max(time.perf_counter() - t0, 1e-6)
    steps = int(max(200, 20 * np.ceil(max(2.0, budget_s) / per20)))
    steps = int(min(steps, 20 * max(10, np.ceil(4.0 * max(2.0, budget_s) / per20))))
    t0 = time.perf_counter()
    loop(sysm, steps)
    dt_s = time.perf_counter() - t0
    val = cfg.n * steps / dt_s
    set_threads(1)
    s1 = max(20, (steps // 4) // 20 * 20)
    t0 = time.perf_counter()
    loop(sysm, s1)
    val1 = cfg.n * s1 / (time.perf_counter() - t0)
    set_threads(threads)
    sysm.close()
    full = {"2": "the full config (10 000 particles, 7.2 GB dense spring table)",
            "3": "the full config (100 000 particles; the unmodified reference would need 2.3 TB of dense tables -> oracle port)",
            "4": f"{args.sample} of {args.copies} aggregates; the full size needs 232 B * N^2 = 232 TB of dense tables",
            "4c": f"{args.sample} of {args.copies} aggregates; the full size needs 232 B * N^2 = 232 TB of dense tables",
            "5": "a 125 000-particle packing at the same volume fraction (oracle port; the reference's 72 B * N^2 table would need 1.1 TB)",
            "5d": "a 50^3 block of the same lattice (oracle port)"}[args.config]
    return {"value": round(val, 1), "unit": METRIC, "cores": int(threads), "kind": kind, "one_thread_value": round(val1, 1),
            "particles": int(cfg.n), "steps": int(steps), "same_config": args.config in ("2", "3"),
            "march": "x86-64-v3 (prebuilt where /root/reference exists; the GPU box has no reference checkout to rebuild with -march=native)",
            "sample": f"{workload_line(desc)} -- {full}; {steps} steps after 40 warm-up steps, constructor excluded",
            "seconds": round(dt_s, 3)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    base = run_reference_sample(args, budget_s=max(2.0, args.cpu_seconds))
    scn_desc = base["sample"]
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": METRIC,
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": base["steps"], "warmup": 40,
        "steps_requested": args.steps, "ms_per_step": round(base["seconds"] / base["steps"] * 1e3, 4), "higher_is_better": True,
        "scaling": "weak" if args.config in ("5", "5d") else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": scn_desc, "config": args.config, "particles": base["particles"], "same_config_as_gpu_arm": base["same_config"]},
        "cpu_baseline": base,
        "e2e": {"value": base["value"], "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="4", choices=["2", "3", "4", "4c", "5", "5d"], help="BASELINE.json config (see the module docstring)")
    ap.add_argument("--copies", type=int, default=6667, help="configs 4 / 4c: aggregates in the workload (6667 -> 1 000 050 particles)")
    ap.add_argument("--sample", type=int, default=34, help="aggregates in the CPU-reference sample (34 -> 5100 particles)")
    ap.add_argument("--variant", type=int, default=None, help="force kernel variant (0 simple, 1 tiled)")
    ap.add_argument("--math", default="fast", choices=["fast", "exact"],
                    help="arithmetic of the step kernel: fast = LG_MATH_FAST (forces within 1e-10 of the reference, sets "
                         "bit-exact; tests/test_gpu_fast_math.py), exact = bit-identical to the reference")
    ap.add_argument("--cpu-seconds", type=float, default=10.0, help="CPU work of the reference sample (seconds; at least 200 steps and 2 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--reorder", default="auto", choices=["auto", "on", "off"],
                    help="internal spatial renumbering (lg_set_reorder); auto = on for the configs whose input order has no locality (2, 3, 5, 5d) in fast math")
    ap.add_argument("--no-autotune", action="store_true", help="keep the deterministic tile-shape rule instead of measuring the shapes")
    ap.add_argument("--kernel-only", action="store_true", help="tuning runs: device-resident throughput + roofline only")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
