#!/usr/bin/env python
"""bench.py — SWE cell-updates/s of the fused fp64 flux + source + Euler stage on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (config C of BASELINE.json): synthetic unstructured triangle floodplain, 2236 x 2236 quads split
into 2 triangles (10.0 M cells) PER GPU, varying bathymetry, Manning n, spatial rain, reflecting walls, semi-implicit
friction, forward Euler.  N > 1: weak scaling over a px x py block decomposition with one ghost layer; the per-step halo
is pushed into the neighbours' ghost slots over NVLink by the compute kernel itself (or NCCL send/recv with --halo nccl).
`--strong`: the SAME px*nx x py*ny mesh cut into N blocks (config E's strong-scaling axis).  A "step" is one fused kernel
pass over every owned cell.  One JSON line on stdout (rank 0).

`checks.parity`: before the timed region the first --parity-steps steps of THIS workload (same mesh, same kernel, same halo
transport, every rank) are compared with the CPU oracle — the reference's own swe_petsc.c operator (oracle/_ref), one per
rank with the ghost exchange of DMGlobalToLocal between them (oracle/partitioned.py).  A failed check exits non-zero.

`--impl reference`: the reference's CPU operator on all host cores, on the same mesh and inputs: rank 0's block of the
workload cut into one sub-block per core with a memcpy halo between them (for N = 1 that is the whole workload).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "SWE cell-updates/s (fp64 RHS+stage)"
UNIT = "cell-updates/s"
BYTES_PER_CELL_TRI = 148  # algorithmic bytes per triangle cell-update, SURVEY §8d / DESIGN.md
FALLBACK_HBM_GBS = 6650.0
PARITY_REL_TOL = 1e-11  # BASELINE.json north_star, floor 1e-3 * scale (tests/util.py); the 1e-6-floor figure is reported beside it
PARITY_NORMWISE_TOL = 1e-13


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=1000)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--nx", type=int, default=2236, help="quads per side of one GPU's block (cells = 2*nx*ny)")
    p.add_argument("--ny", type=int, default=0)
    p.add_argument("--kernel", default=os.environ.get("RDYB_KERNEL", "auto"), choices=["auto", "gather", "tiled", "async", "fixed"])
    p.add_argument("--tile", type=int, default=None, help="cells per tile (default: RDYB_TILE, else 224 = K4's shape, 192 with --second-order = K4M's shape)")
    p.add_argument("--no-courant", action="store_true", help="skip the per-edge Courant diagnostic (the reference always computes it)")
    p.add_argument("--e2e-steps", type=int, default=10)
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="target duration of the cpu_baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--halo", default="p2p", choices=["p2p", "nccl"], help="N>1: peer-memory push over NVLink (CUDA IPC) or NCCL send/recv")
    p.add_argument("--eta0", type=float, default=None, help="initial lake level (default 18 m: 84 %% of the cells wet at N=1; 60 m = fully wet)")
    p.add_argument("--hr", action="store_true", help="config D: hydrostatic reconstruction + dam column")
    p.add_argument("--second-order", action="store_true",
                   help="secondary line: numerics.second_order (MUSCL, minmod) on a fully wet config C (eta0 60 m unless --eta0; the scheme has no dry-cell treatment)")
    p.add_argument("--rk4", action="store_true", help="secondary line: classical RK4 (TSRK4) = 4 RHS evaluations + stage AXPYs per step instead of the fused Euler step")
    p.add_argument("--strong", action="store_true", help="N>1: strong scaling - --nx/--ny name the WHOLE mesh, cut into N blocks")
    p.add_argument("--parity-steps", type=int, default=5, help="steps of this workload compared with the CPU oracle before the timed region (0 = skip)")
    p.add_argument("--max-ctas", type=int, default=0, help="experiments: cap the persistent grid")
    return p.parse_args()


def block_grid(n):
    """px x py block decomposition for n ranks."""
    if n in (1, 2, 4, 8, 16, 32):
        return {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2), 16: (4, 4), 32: (8, 4)}[n]
    px = int(np.floor(np.sqrt(n)))
    while n % px:
        px -= 1
    return (n // px, px)


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _sha16(path):
    try:
        with open(path, "rb") as f:
            return hashlib.sha256(f.read()).hexdigest()[:16]
    except Exception:
        return None


def recorded_traffic(kernel_name):
    """DRAM bytes per cell of the dominant kernel from the committed `ncu --set full` capture (profiles/traffic.json): an entry is
    {"bytes_per_cell", "git", "sources": {file: sha16 at capture time}}.  Returns (bytes_per_cell or None, note): the figure is
    withheld when a kernel source changed after the capture (a stale number would survive any kernel edit otherwise)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            ent = json.load(f).get(kernel_name)
    except Exception:
        return None, "no profiles/traffic.json"
    if ent is None:
        return None, f"no ncu capture recorded for {kernel_name}"
    if not isinstance(ent, dict):
        return None, "unstamped entry (no source hashes): not reported"
    for rel, sha in ent.get("sources", {}).items():
        if _sha16(os.path.join(ROOT, rel)) != sha:
            return None, f"{rel} changed after the capture at {ent.get('git')}: re-capture needed"
    return float(ent["bytes_per_cell"]), f"ncu --set full capture at {ent.get('git')} ({ent.get('file', 'profiles/')})"


# --------------------------------------------------------------------------------------------------------------
# the workload: one definition for both arms
# --------------------------------------------------------------------------------------------------------------
def workload(args, world):
    """(px, py, whole-mesh nx, ny, eta0, kwargs of cases.floodplain) of this run."""
    ny_arg = args.ny or args.nx
    px, py = block_grid(world)
    if args.strong:
        nx, ny = args.nx, ny_arg
    else:
        nx, ny = args.nx * px, ny_arg * py
    if args.hr:
        eta0 = 12.0 if args.eta0 is None else args.eta0
        kw = dict(nx=nx, ny=ny, eta0=eta0, hr=True, dam_column=(2000.0, 28.0))
    else:
        eta0 = (60.0 if args.second_order else 18.0) if args.eta0 is None else args.eta0
        kw = dict(nx=nx, ny=ny, eta0=eta0)
    return px, py, nx, ny, eta0, kw


def workload_dt(args, eta0):
    """One time step for every rank and both arms: the Courant bound of cases.floodplain with the deepest water the domain can hold
    (the bed's lowest point is about -5 zamp m; the dam column of config D is 28 m high)."""
    zamp = 1.5 if args.hr else 1.0
    hmax = max(eta0 + 5.0 * zamp, 28.0 + 5.0 * zamp if args.hr else 0.0, 1.0)
    return float(0.25 * 0.2 * 10.0 / np.sqrt(9.806 * hmax))


def config_dict(args, world, cells_total, wet_fraction, dt):
    """The `config` object: the same keys and values in both arms (the driver compares them)."""
    px, py, nx, ny, eta0, _ = workload(args, world)
    name = "config D: HR floodplain with a dam column" if args.hr else "config C: synthetic unstructured triangle floodplain"
    scheme = "second order (MUSCL, minmod)" if args.second_order else "first order"
    return {
        "workload": f"{name}, {nx}x{ny}x2 = {2 * nx * ny} triangles over {world} GPU(s) ({px}x{py} blocks, "
                    f"{'strong' if args.strong else 'weak'} scaling), bathymetry + Manning n + rain, reflecting walls, "
                    f"{'hydrostatic-reconstruction' if args.hr else 'Roe'} flux, semi-implicit friction, {scheme}, "
                    f"{'classical RK4' if args.rk4 else 'forward Euler'}, eta0 = {eta0:g} m",
        "cells_total": int(cells_total), "wet_fraction": round(float(wet_fraction), 4), "dt": dt,
        "l2": "inputs larger than L2 (state 240 MB + mesh > 1 GB per GPU vs 126 MB L2); no flush needed",
    }


# --------------------------------------------------------------------------------------------------------------
# clocks: sampled with NVML during the timed region
# --------------------------------------------------------------------------------------------------------------
class ClockSampler:
    def __init__(self, device_index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nv = pynvml
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[device_index]) if visible and visible.split(",")[device_index].isdigit() else device_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None

    def _sample(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        try:
            self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
            mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for k, bit in names.items():
                if mask & bit:
                    self.reasons.add(k)
        except Exception:
            pass

    def _loop(self):
        while not self._stop.is_set():
            self._sample()
            self._stop.wait(0.002)

    def start(self):
        if self.nv:
            self._thread = threading.Thread(target=self._loop, daemon=True)
            self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join()
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


# --------------------------------------------------------------------------------------------------------------
# CPU arm: the reference's own CPU operator (oracle/_ref, else the bit-identical port) on the host cores
# --------------------------------------------------------------------------------------------------------------
def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


class CpuArm:
    """Rank 0's block of the bench workload (at N = 1: the whole mesh) on P host threads: the block is cut into sx x sy sub-blocks
    by the same per-partition generator, one reference operator per sub-block, ghosts copied between them before every RHS
    (DMGlobalToLocal as a memcpy; no PetscSF/MPI cost: this flatters the CPU).  Ghosts whose owner lies outside the sample (N > 1
    only) stay frozen at their initial value.  ctypes releases the GIL, so the threads run in parallel."""

    def __init__(self, args, world, cores=None):
        from oracle import binding
        from oracle.partitioned import InProcessParts, Part
        from rdycore_b200 import cases

        self.kind = binding.best_kind()
        self.cores = cores or host_cores()
        px, py, nx, ny, eta0, kw = workload(args, world)
        sx, sy = block_grid(self.cores)
        self.dt = workload_dt(args, eta0)
        self.blocks = (sx, sy)
        parts = []
        # sub-block (i, j) of GPU block (0, 0) = block (i, j) of the (px*sx) x (py*sy) decomposition of the whole mesh
        for j in range(sy):
            for i in range(sx):
                case = cases.floodplain(px=px * sx, py=py * sy, rank=j * (px * sx) + i, dt=self.dt, **kw)
                parts.append(Part(case, self.kind))
        self.job = InProcessParts(parts)
        self.cells_per_step = self.job.cells_per_step
        self.wet_fraction = float(sum((p.case.u0[:, 0] > 0).sum() for p in parts)) / self.cells_per_step
        self.desc0 = (f"rank 0's block of the workload ({self.cells_per_step} cells) as {sx}x{sy} sub-blocks on {self.cores} host threads, "
                      f"memcpy halo, oracle kind '{self.kind}' (the reference's src/swe/swe_petsc.c operator; no MPI/PetscSF cost)")

    def run(self, nsteps):
        return self.job.step_euler(self.dt, nsteps)

    def sample_desc(self, nsteps):
        return f"{nsteps} Euler steps of {self.desc0}"


def cpu_baseline(args, world, seconds):
    """P = all cores and P = 1 on the same mesh and inputs."""
    arm = CpuArm(args, world)
    t1 = arm.run(2)  # warm-up + calibration
    nsteps = max(3, int(seconds / max(t1 / 2, 1e-6)))
    t = arm.run(nsteps)
    out = {"value": arm.cells_per_step * nsteps / t, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample_desc(nsteps)}
    # P = 1: one sub-block, one thread, its ghosts frozen (a bounded sample of the same mesh)
    from oracle.partitioned import InProcessParts

    one = InProcessParts(arm.job.parts[:1])
    one.step_euler(arm.dt, 1)
    n1 = max(2, int(0.25 * seconds / max(t1 / 2, 1e-6)))
    t = one.step_euler(arm.dt, n1)
    out["p1"] = {"value": one.cells_per_step * n1 / t, "unit": UNIT, "cores": 1, "sample": f"{n1} Euler steps of sub-block 0 ({one.cells_per_step} cells), ghosts frozen"}
    return out


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return
    world = max(world, args.gpus) if world > 1 else args.gpus
    arm = CpuArm(args, world)
    arm.run(max(1, min(args.warmup, 5)))
    t = arm.run(args.steps)
    value = arm.cells_per_step * args.steps / t
    px, py, nx, ny, eta0, _ = workload(args, world)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": config_dict(args, world, 2 * nx * ny, arm.wet_fraction, arm.dt),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample_desc(args.steps)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "detail": {"sample_cells": arm.cells_per_step, "sub_blocks": list(arm.blocks),
                   "note": "throughput of the bounded sample; at N > 1 the sample is one GPU's block of the workload"},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------------
# parity of the bench workload itself
# --------------------------------------------------------------------------------------------------------------
def parity_check(args, op, case, dt, world, rank, dist, step_fn):
    """The first --parity-steps steps of THIS run against the CPU oracle on the same mesh and inputs."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from util import state_error  # the parity metric of the test suite

    from oracle.partitioned import DistributedPart, Part

    t0 = time.perf_counter()
    steps = args.parity_steps
    step_fn(steps - 1)
    op.reset_diagnostics()  # the oracle's Courant number is the one of its last RHS evaluation
    step_fn(1)
    op.synchronize()
    u = op.get_state()
    courant = op.courant().max_courant_num
    if world == 1:
        ref = Part(case)
        for _ in range(steps):
            ref.step(dt)
        u_ref, c_ref, kind = ref.u_global(), ref.oracle.courant(), ref.kind
    else:
        ref = DistributedPart(case, dist, rank)
        ref.step_euler(dt, steps)
        u_ref, c_ref, kind = ref.u_global(), ref.courant(), ref.part.kind
    # the scales of the relative error are those of the WHOLE mesh: a rank whose block is a lake at rest has momenta of 1e-9, and
    # round-off measured against them is not a parity figure
    local_scales = (float(np.abs(u_ref[:, 0]).max()), float(np.sqrt(u_ref[:, 1] ** 2 + u_ref[:, 2] ** 2).max()))
    if world > 1:
        alls = [None] * world
        dist.all_gather_object(alls, local_scales)
    else:
        alls = [local_scales]
    scales = (max(a[0] for a in alls), max(a[1] for a in alls))
    rel3, abs_dry, mism, normwise = state_error(u, u_ref, rel_floor=1e-3, scales=scales)
    rel6, _, _, _ = state_error(u, u_ref, rel_floor=1e-6, scales=scales)
    moved = float(np.abs(u_ref[:, 1:]).max())
    mine = [rel3, rel6, normwise, abs_dry, float(mism), moved, courant, c_ref, 0.0 if np.isfinite(u).all() else 1.0]
    if world > 1:
        allv = [None] * world
        dist.all_gather_object(allv, mine)
    else:
        allv = [mine]
    a = np.array(allv)
    cg, cr = a[:, 6].max(), a[:, 7].max()
    out = {
        "max_rel": float(a[:, 0].max()), "max_rel_floor_1e-6": float(a[:, 1].max()), "normwise": float(a[:, 2].max()),
        "max_abs_near_dry": float(a[:, 3].max()), "wetdry_mismatch": int(a[:, 4].sum()), "cells": int(2 * workload(args, world)[2] * workload(args, world)[3]),
        "steps": steps, "ranks": world, "oracle": kind, "max_momentum_ref": float(a[:, 5].max()),
        "courant_gpu": float(cg), "courant_ref": float(cr), "courant_rel_diff": float(abs(cg - cr) / max(cr, 1e-300)),
        "tolerance": {"max_rel": PARITY_REL_TOL, "floor": "1e-3 * scale (tests/util.py); max_rel_floor_1e-6 is SURVEY 8c's figure, reported", "normwise": PARITY_NORMWISE_TOL},
        "seconds": round(time.perf_counter() - t0, 1),
    }
    out["ok"] = bool(a[:, 8].max() == 0.0 and out["max_rel"] <= PARITY_REL_TOL and out["normwise"] <= PARITY_NORMWISE_TOL and out["max_abs_near_dry"] <= 1e-12
                     and out["courant_rel_diff"] <= 1e-12 and out["max_momentum_ref"] > 0.0)
    return out


# --------------------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    from rdycore_b200 import cases
    from rdycore_b200._lib import lib
    from rdycore_b200.operator import GATHER, TILED, TILED_ASYNC, TILED_FIXED, Operator, kernel_launch_count

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched with torch.distributed.run --nproc-per-node N")
        args.gpus = world
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the SWE path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    if args.tile is None:
        args.tile = int(os.environ.get("RDYB_TILE", "192" if args.second_order else "224"))
    px, py, nx, ny, eta0, kw = workload(args, world)
    dt = workload_dt(args, eta0)
    t_setup = time.perf_counter()
    case = cases.floodplain(px=px, py=py, rank=rank, dt=dt, **kw)
    if args.second_order:
        case.config.reconstruction = cases.MUSCL_MINMOD
    halo = case.mesh.halo() if world > 1 else None
    op = Operator.from_case(case, device=local_rank, halo=halo)
    kernel = {"auto": TILED_FIXED, "gather": GATHER, "tiled": TILED, "async": TILED_ASYNC, "fixed": TILED_FIXED}[args.kernel]
    kernel_name = {GATHER: "gather", TILED: "tiled", TILED_ASYNC: "tiled_async", TILED_FIXED: "tiled_fixed"}[kernel]
    op.set_kernel(kernel, args.tile)
    if args.max_ctas:
        op.set_max_ctas(args.max_ctas)
    if args.second_order:
        # the second-order path runs muscl_gradient_kernel + K2M on any tiled layout, + K1<RECON> with --kernel gather
        kernel_name = "gather_muscl" if kernel == GATHER else "tiled_muscl"
    op.set_options(courant=not args.no_courant)
    if world > 1:
        uid = [Operator.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        op.init_comm(world, rank, uid[0])
        if args.halo == "p2p":
            blobs = [None] * world
            dist.all_gather_object(blobs, op.halo_export(rank))
            op.halo_import(world, rank, blobs)
    t_setup = time.perf_counter() - t_setup
    n_owned = op.num_owned
    stages = 4 if args.rk4 else 1

    def step_fn(n):
        if args.rk4:
            op.step_rk4(dt, n)
        else:
            op.step_euler(dt, n)

    def barrier():
        op.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def gather_ranks(x):
        if world == 1:
            return [x]
        out = [None] * world
        dist.all_gather_object(out, x)
        return out

    total_cells = float(sum(gather_ranks(n_owned)))
    wet_fraction = float(sum(gather_ranks(int((case.u0[:, 0] > 0).sum())))) / total_cells

    # ---- parity of this very workload (mesh, kernel, halo transport) against the CPU oracle, before anything is timed
    parity = None
    if args.parity_steps > 0:
        parity = parity_check(args, op, case, dt, world, rank, dist if world > 1 else None, step_fn)

    # ---- device-resident timing: W warm-up steps, then exactly K steps between CUDA events on the operator's stream.
    # Everything slow on the host (NVML init) happens BEFORE the barrier; with the peer-memory halo the start event is recorded on
    # the device after every neighbour has delivered its priming push (rdyb_operator_step_euler), so host launch skew between the
    # ranks is not inside the timed region.
    step_fn(args.warmup)
    sampler = ClockSampler(local_rank)
    op.wait_cycles()  # reset the in-kernel wait counters
    barrier()
    sampler.start()
    n0 = kernel_launch_count()
    t_host = time.perf_counter()
    step_fn(args.steps)
    ms_per_step, _ = op.last_kernel_ms()  # CUDA events on the operator's stream around the K steps; event-synchronises
    t_host = time.perf_counter() - t_host
    kinfo = op.kernel_info()
    launches = kernel_launch_count() - n0
    wait_sum, wait_max, wait_n = op.wait_cycles()
    clocks = sampler.stop()
    barrier()
    per_rank_ms = gather_ranks(ms_per_step)
    per_rank_wait = gather_ranks((wait_sum, wait_max, wait_n))
    per_rank_clocks = gather_ranks(clocks)
    ms_per_step = max(per_rank_ms)
    value = total_cells * stages / (ms_per_step * 1e-3)
    state = op.get_state()
    finite = bool(np.isfinite(state).all())
    courant = op.courant().max_courant_num

    # ---- e2e: the reference-facing call (PetscOperator apply shape) with HOST Vec arrays, copies inside the timed region
    nl = op.num_local
    import ctypes as C

    pu, pf = C.c_void_p(), C.c_void_p()
    lib().rdyb_host_alloc(C.c_size_t(8 * 3 * nl), C.byref(pu))
    lib().rdyb_host_alloc(C.c_size_t(8 * 3 * n_owned), C.byref(pf))
    u_host = np.ctypeslib.as_array(C.cast(pu, C.POINTER(C.c_double)), shape=(3 * nl,))
    f_host = np.ctypeslib.as_array(C.cast(pf, C.POINTER(C.c_double)), shape=(3 * n_owned,))
    u_host[:] = 0.0
    u_host.reshape(-1, 3)[: n_owned] = state  # local order = owned first for the generated meshes; ghosts stay as uploaded
    for _ in range(3):
        op.apply_host_ptr(dt, pu.value, pf.value)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        op.apply_host_ptr(dt, pu.value, pf.value)  # returns after the D2H copy of F completed
    barrier()
    e2e_s = max(gather_ranks((time.perf_counter() - t0) / args.e2e_steps))
    e2e = {"value": total_cells / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(sum(gather_ranks(24 * nl))),
           "d2h_bytes_per_step": int(sum(gather_ranks(24 * n_owned))), "ms_per_step": 1e3 * e2e_s,
           "call": "rdyb_operator_apply_host (PetscOperator apply shape: host u_local AoS in, host F AoS out, pinned)"}
    f_finite = bool(np.isfinite(f_host).all())
    lib().rdyb_host_free(pu)
    lib().rdyb_host_free(pf)

    ok = True
    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_per_cell = 140 if args.hr else BYTES_PER_CELL_TRI
        fused_gradients = args.second_order and "in-kernel gradients" in kinfo  # K4M: no gradient array, no gradient pass
        if args.second_order:  # + per-slot LS weights (3 slots x 2 doubles) + face offsets (K1: 3 slots x 4; tiles: 1.5 edges x 4) + gradients written and read (48 + 48) unless fused
            bytes_per_cell += (0 if fused_gradients else 96) + 48 + (96 if kernel == GATHER else 48)
        bytes_per_step_cell = bytes_per_cell
        if args.rk4:  # 4 RHS evaluations + 3 stage AXPYs (read U, K; write Y: 72 B) + the combine (read U, 4 K; write U: 144 B), BASELINE.md §3
            bytes_per_step_cell = 4 * bytes_per_cell + 3 * 72 + 144
        achieved = bytes_per_step_cell * n_owned / (ms_per_step * 1e-3) / 1e9  # this rank's kernel(s): per-GPU figure
        traffic, traffic_note = recorded_traffic(kernel_name + ("_hr" if args.hr else ""))
        if args.rk4 or world > 1 and args.strong:
            traffic, traffic_note = None, "captured for the fused Euler step on the 10 M-cell block only"
        sm_clock = clocks.get("sm_mhz") or 1965.0
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong" if args.strong else "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": config_dict(args, world, total_cells, wet_fraction, dt),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (traffic * n_owned if traffic else None), "traffic_note": traffic_note, "peak_source": peak_src,
                         "algorithmic_bytes_per_cell": bytes_per_cell, "algorithmic_bytes_per_cell_per_step": bytes_per_step_cell,
                         "kernel": ("muscl_gradient_kernel + " + ("swe_gather_kernel<RECON>" if kernel == GATHER else "swe_tiled_muscl_async_kernel (K3M)")) if args.second_order and not fused_gradients else "swe_tiled_muscl_fixed_kernel (K4M, gradients in-kernel)" if args.second_order else "swe_%s_kernel" % kernel_name,
                         "note": "per-GPU: algorithmic bytes of one step on rank 0 / CUDA-event time per step on the launching stream (max over ranks)"},
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "checks": {"state_finite": finite, "f_finite": f_finite, "max_courant": courant, "parity": parity},
            "detail": {
                "decomposition": (f"{px}x{py} blocks, 1 ghost layer, per-step halo: " + ("border cells stored into the neighbours' ghost slots over NVLink by the compute kernel (CUDA IPC peer memory)" if args.halo == "p2p" else "NCCL send/recv")) if world > 1 else "single GPU",
                "kernel_info": kinfo, "kernel": kernel_name, "tile_cells": args.tile if kernel != GATHER else None,
                "courant_diagnostic": not args.no_courant, "setup_s": round(t_setup, 2), "cells_rank0": int(n_owned),
                "per_rank_ms_per_step": [round(float(x), 5) for x in per_rank_ms],
                "per_rank_mean_wait_us": [round(w[0] / sm_clock / max(w[2], 1), 3) for w in per_rank_wait],
                "per_rank_waiting_ctas_per_step": [round(w[2] / max(args.steps, 1), 1) for w in per_rank_wait],
                "per_rank_longest_wait_us": [round(w[1] / sm_clock, 1) for w in per_rank_wait],
                "per_rank_sm_mhz": [c.get("sm_mhz") for c in per_rank_clocks],
                "host_enqueue_ms_per_step": round(1e3 * t_host / max(args.steps, 1), 4),
                "wait_note": "in-kernel neighbour wait: the producer warp of every CTA that owns a frontier tile waits once per step for its neighbours' border cells (concurrently with the other CTAs); mean and longest single wait in us (clock64 ticks / SM MHz)",
            },
        }
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args, world, args.cpu_seconds)
        print(json.dumps(line), flush=True)
        if parity is not None and not parity["ok"]:
            sys.stderr.write("bench.py: PARITY CHECK FAILED: %s\n" % json.dumps(parity))
            ok = False
    op.close()
    if world > 1:
        flag = [ok]
        dist.broadcast_object_list(flag, src=0)
        ok = flag[0]
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        sys.exit(3)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
