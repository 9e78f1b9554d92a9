#!/usr/bin/env python
"""bench.py -- particle-timesteps/s of the DEM hot path on BASELINE.json config 4.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--copies C]

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
from ctypes import byref, c_double as C_double, c_int64 as C_int64

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from libgran_b200 import scenarios as S  # noqa: E402

METRIC = "particle-timesteps/s"
# dram__bytes_read.sum + dram__bytes_write.sum per launch of the fused step kernel on this workload, from the committed
# `ncu --set full` captures (profiles/r01d_step_tiled_ncu_raw.csv: 482.0 MB + 301.7 MB; exact mode r01b: 482 + 301 MB)
TRAFFIC_NCU = {"fast": 806.3e6, "exact": 783.0e6}  # fast: step_rows_kernel, profiles/r01f_step_rows_ncu_raw.csv (491.9 + 314.4 MB)
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


def build_workload(copies: int):
    scn = S.multi_aggregate_collision(copies)
    return scn


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

    scn = build_workload(args.copies)
    cfg = scn.cfg
    n = cfg.n
    if world > 1:
        return run_slab(args, scn, world, rank, local_rank)

    uv = numpy_unit_vectors(n + 1024)
    mode = 1 if (args.math == "fast" and args.variant in (None, 1)) else 0
    sysm = from_config(cfg, *scn.state, unit_vectors=uv, device=local_rank, variant=args.variant, math_mode=mode)
    stream = torch.cuda.ExternalStream(sysm.stream_handle(), device=torch.device("cuda", local_rank))

    def timed_run(k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        with torch.cuda.stream(stream):
            e0.record()
            sysm.run(cfg.dt, k, REBUILD_EVERY)
            e1.record()
        sysm.synchronize()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    # ---- device-resident throughput ------------------------------------------------------------
    sysm.run(cfg.dt, effective_warmup(args), REBUILD_EVERY)
    sysm.synchronize()
    c0 = sysm.counters()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ms = timed_run(args.steps)
    clocks = sampler.stop()
    c1 = sysm.counters()
    value = n * args.steps / (ms * 1e-3)
    launches = c1["launches"] - c0["launches"]

    # ---- roofline of the dominant kernel (pair force), measured live with events ----------------
    sysm.set_profiling(True)
    prof_steps = min(args.steps, 40)
    sysm.run(cfg.dt, prof_steps, REBUILD_EVERY)
    force_ms, force_launches = sysm.get_profile()
    sysm.set_profiling(False)
    cnt = sysm.counters()
    step_bytes, force_bytes = algorithmic_bytes(n, cnt["pairs"], cnt["contacts"], cnt["bonds"])
    peak, peak_src = measured_peaks()
    force_avg_ms = force_ms / max(force_launches, 1)
    # the production launch is the FUSED step kernel (force + Verlet + connector update): it moves the whole B_step;
    # the validation variant's force-only kernel moves force_bytes
    fused = args.variant in (None, 1)
    launch_bytes = step_bytes if fused else force_bytes
    achieved = launch_bytes / (force_avg_ms * 1e-3) / 1e9
    rows_kernel = fused and mode == 1 and cnt["pairs"] <= 24 * n and os.environ.get("LGB200_ROWS_KERNEL", "1") != "0"
    kname = ("step_rows_kernel" if rows_kernel else "step_tiled_kernel") + " (fused pair force + Verlet + connectors)"
    roofline = {"bound": "hbm", "kernel": kname if fused else "pair_force",
                "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                "frac": round(achieved / peak, 4), "traffic": TRAFFIC_NCU.get(args.math if fused else "variant0"), "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(launch_bytes), "avg_launch_ms": round(force_avg_ms, 5),
                "launches_timed": int(force_launches),
                "share_of_step": round(force_avg_ms / (ms / args.steps), 3),
                "whole_step": {"algorithmic_bytes_per_step": int(step_bytes),
                               "achieved_gbs": round(step_bytes * args.steps / (ms * 1e-3) / 1e9, 1),
                               "frac": round(step_bytes * args.steps / (ms * 1e-3) / 1e9 / peak, 4),
                               "frac_of_nominal_8TBs": round(step_bytes * args.steps / (ms * 1e-3) / 8e12, 4)}}

    if args.kernel_only:
        if rank == 0:
            print(json.dumps({"math": args.math, "value": round(value, 1), "ms_per_step": round(ms / args.steps, 5), "roofline": roofline,
                              "gpu_launches": int(launches), "clocks": clocks, "lib": os.environ.get("LGB200_LIB", "default"), "tag": os.environ.get("LGB200_TAG")}))
        return

    # ---- end to end through the host-facing API with host buffers ---------------------------------
    pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in scn.state]
    host_np = [t.numpy() for t in pinned]
    outs = [torch.empty((n, 3), dtype=torch.float64).pin_memory() for _ in range(4)]
    out_np = [t.numpy() for t in outs]
    e2e_sys = from_config(cfg, *scn.state, unit_vectors=uv, device=local_rank, variant=args.variant, math_mode=mode)  # model tables built once (as a ctor would)
    e2e_steps = args.steps
    # untimed priming with dt = 0: sizes the neighbor / scratch buffers for exactly this state and leaves positions,
    # velocities and bridge connectors where the constructor put them (x += v * 0, c += 0)
    e2e_sys.run(0.0, effective_warmup(args), REBUILD_EVERY)
    e2e_sys.synchronize()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e2e_sys.set_state(*host_np)  # H2D of x0, v0, theta0, omega0 (re-arms the first-step half kick)
    ke = 0.0
    for k in range(e2e_steps):
        if k % REBUILD_EVERY == 0:
            e2e_sys.update_neighbor_list()
        e2e_sys.do_step(cfg.dt)
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
    e2e_sys.close()

    line = {
        "metric": METRIC, "value": round(value, 1), "unit": METRIC, "n_gpus": world, "steps": args.steps,
        "warmup": effective_warmup(args), "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"cfg4 multi-aggregate collision, {n} particles ({args.copies} x aggregate.h), sinter+contact+Hamaker, "
                               f"r_verlet=5r, rebuild every {REBUILD_EVERY} steps", "particles": n, "pairs": cnt["pairs"],
                   "bonds_ordered": cnt["bonds"], "contacts": cnt["contacts"], "l2_policy": "working set (>400 MB) larger than L2",
                   "kernel_variant": args.variant, "math_mode": "fast (<=1e-10 rel. forces, bit-exact sets)" if mode else "exact (bit-identical to the reference)"},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline,
    }
    if rank == 0 and not args.no_cpu_baseline:
        line["cpu_baseline"] = run_reference_sample(args, steps=args.cpu_steps, warmup=5)
    if rank == 0:
        print(json.dumps(line))


def run_slab(args, scn, world, rank, local_rank):
    """N > 1: the same workload slab-decomposed along x, one rank per GPU (libgran_b200/slab.py).  Strong scaling:
    the particle count is fixed, `value` = all particles * steps / max-over-ranks device time."""
    import torch
    import torch.distributed as dist

    from libgran_b200.slab import SlabSystem

    cfg = scn.cfg
    n = cfg.n
    dev = torch.device("cuda", local_rank)
    uv = numpy_unit_vectors(n + 1024)
    mode = 1 if args.math == "fast" else 0
    sysm = SlabSystem(cfg, *scn.state, unit_vectors=uv, device=local_rank, math_mode=mode)
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
    roofline = {"bound": "hbm", "kernel": ("step_rows_kernel" if mode else "step_tiled_kernel") + " (fused pair force + Verlet + connectors), this rank", "achieved": round(ach, 1),
                "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": int(step_bytes_local), "avg_launch_ms": round(force_avg_ms, 5),
                "slowest_rank_launch_ms": round(slow, 5), "launches_timed": int(fl.value),
                "share_of_step": round(force_avg_ms / (ms / args.steps), 3)}

    # end to end: host buffers in, host buffers out, through the slab driver
    # (a fresh system, primed with dt = 0 like the single-GPU arm: same buffers, untouched state and connectors)
    sysm.close()
    sysm = SlabSystem(cfg, *scn.state, unit_vectors=uv, device=local_rank, math_mode=mode)
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
    nloc = torch.tensor([sysm.n_local, own], dtype=torch.int64, device=dev)
    dist.all_reduce(nloc)
    e2e = {"value": round(n * e2e_steps / float(tt), 1), "unit": METRIC,
           "h2d_bytes_per_step": round(4 * 24 * n * world / e2e_steps, 1),
           "d2h_bytes_per_step": round((4 * 24 * int(nloc[1]) + (e2e_steps // 10) * 12 * 8 * world) / e2e_steps, 1),
           "steps": e2e_steps, "final_ke": ke,
           "what": "per rank: set_state_host (H2D, pinned) + K/10 x [run(10 steps: step + halo exchange, rebuild/20) + KE "
                   "all-reduce + read-back] + owned x,v,omega,theta (D2H)"}
    del outs
    line = {
        "metric": METRIC, "value": round(value, 1), "unit": METRIC, "n_gpus": world, "steps": args.steps, "warmup": warm,
        "ms_per_step": round(ms / args.steps, 5), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"cfg4 multi-aggregate collision, {n} particles ({args.copies} x aggregate.h), sinter+contact+Hamaker, "
                               f"r_verlet=5r, rebuild every {REBUILD_EVERY} steps", "particles": n, "pairs": cnt["pairs"],
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


def run_reference_sample(args, steps, warmup):
    """The reference's own OpenMP implementation on a bounded sample of the workload."""
    os.environ.setdefault("OMP_PROC_BIND", "close")
    from oracle import bindings as ob

    sample = args.sample
    scn = build_workload(sample)
    cfg = scn.cfg
    kind = "reference"
    if ob.ref_available(timing=True):
        threads = ob.load_ref(timing=True).ref_max_threads()
        sysm = ob.RefSystem(cfg, *scn.state, timing=True)
    else:
        kind = "port"
        threads = ob.load_oracle().orc_max_threads()
        sysm = ob.OracleSystem(cfg, *scn.state)
    sysm.run(cfg.dt, warmup, REBUILD_EVERY)
    t0 = time.perf_counter()
    sysm.run(cfg.dt, steps, REBUILD_EVERY)
    dt_s = time.perf_counter() - t0
    val = cfg.n * steps / dt_s
    sysm.close()
    return {"value": round(val, 1), "unit": METRIC, "cores": int(threads), "kind": kind,
            "sample": f"{sample} of {args.copies} aggregates ({cfg.n} particles), {steps} steps after {warmup} warm-up, "
                      f"constructor excluded; full size infeasible (232 B * N^2 dense tables)",
            "seconds": round(dt_s, 3)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    scn_n = 150 * args.copies
    base = run_reference_sample(args, steps=args.steps, warmup=max(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": base["value"], "unit": METRIC,
        "n_gpus": int(os.environ.get("WORLD_SIZE", "1")), "steps": args.steps, "warmup": max(args.warmup, 1),
        "ms_per_step": round(base["seconds"] / args.steps * 1e3, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"cfg4 multi-aggregate collision, {scn_n} particles ({args.copies} x aggregate.h), "
                               f"sinter+contact+Hamaker, r_verlet=5r, rebuild every {REBUILD_EVERY} steps",
                   "particles": scn_n},
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
    ap.add_argument("--copies", type=int, default=6667, help="aggregates in the workload (6667 -> 1 000 050 particles)")
    ap.add_argument("--sample", type=int, default=34, help="aggregates in the CPU-reference sample (34 -> 5100 particles)")
    ap.add_argument("--variant", type=int, default=None, help="force kernel variant (0 simple, 1 tiled)")
    ap.add_argument("--math", default="fast", choices=["fast", "exact"],
                    help="arithmetic of the step kernel: fast = LG_MATH_FAST (forces within 1e-10 of the reference, sets "
                         "bit-exact; tests/test_gpu_fast_math.py), exact = bit-identical to the reference")
    ap.add_argument("--cpu-steps", type=int, default=10000, help="time steps of the CPU-reference sample (about 10 s of CPU work)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--kernel-only", action="store_true", help="tuning runs: device-resident throughput + roofline only")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
