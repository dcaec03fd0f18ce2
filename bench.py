#!/usr/bin/env python
"""bench.py — SWE cell-updates/s of the fused fp64 flux + source + Euler stage on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

Workload (config C of BASELINE.json): synthetic unstructured triangle floodplain, 2236 x 2236 quads split
into 2 triangles (10.0 M cells) PER GPU, varying bathymetry, Manning n, spatial rain, reflecting walls, semi-implicit
friction, forward Euler.  N > 1: weak scaling over a px x py block decomposition with one ghost layer and a per-step
halo exchange (NCCL send/recv) overlapped with the interior cells.  A "step" is one fused kernel pass over every owned
cell.  One JSON line on stdout (rank 0).
"""
from __future__ import annotations

import argparse
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


def parse_args():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=1000)
    p.add_argument("--warmup", type=int, default=20)
    p.add_argument("--impl", default="b200", choices=["b200", "reference"])
    p.add_argument("--nx", type=int, default=2236, help="quads per side of one GPU's block (cells = 2*nx*ny)")
    p.add_argument("--ny", type=int, default=0)
    p.add_argument("--kernel", default=os.environ.get("RDYB_KERNEL", "auto"), choices=["auto", "gather", "tiled", "async", "fixed", "warp", "split"])
    p.add_argument("--tile", type=int, default=None, help="cells per tile (default: RDYB_TILE, else 224 = K4's shape, 128 with --second-order = K2M's)")
    p.add_argument("--no-courant", action="store_true", help="skip the per-edge Courant diagnostic (the reference always computes it)")
    p.add_argument("--e2e-steps", type=int, default=10)
    p.add_argument("--cpu-seconds", type=float, default=12.0, help="target duration of the cpu_baseline sample")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--halo", default="p2p", choices=["p2p", "nccl"], help="N>1: peer-memory push over NVLink (CUDA IPC) or NCCL send/recv")
    p.add_argument("--eta0", type=float, default=None, help="initial lake level (default 18 m: ~55 %% of the cells wet; experiments only)")
    p.add_argument("--hr", action="store_true", help="config D: hydrostatic reconstruction + dam column")
    p.add_argument("--second-order", action="store_true",
                   help="secondary line: numerics.second_order (MUSCL, minmod) on a fully wet config C (eta0 60 m unless --eta0; the scheme has no dry-cell treatment)")
    return p.parse_args()


def block_grid(n):
    """px x py block decomposition for n ranks."""
    return {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}.get(n, (n, 1))


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def recorded_traffic(kernel_name):
    """DRAM bytes per cell of the dominant kernel from the committed ncu --set full capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(kernel_name)
    except Exception:
        return None


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

    def _loop(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self._stop.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._stop.wait(0.01)

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
    """One oracle per host core, each on its own block of the same floodplain generator (no halo exchange between
    blocks and no PetscSF/MPI cost: this flatters the CPU).  ctypes releases the GIL, so threads run in parallel."""

    def __init__(self, block_nx, hr=False):
        from oracle import binding
        from rdycore_b200 import cases

        self.kind = binding.best_kind()
        self.cores = host_cores()
        self.block_nx = block_nx
        self.oracles, self.cases = [], []
        for c in range(self.cores):
            case = cases.floodplain(nx=block_nx, ny=block_nx, eta0=8.0 if not hr else 6.0, hr=hr)
            arr, nb = case.boundary_views()
            view = case.mesh.view()
            o = binding.Oracle(self.kind, case.config, view, arr)
            o._keep2 = (arr, view)
            o.array("material_properties")[:] = case.mannings
            o.array("external_sources")[:] = case.sources.ravel()
            o.array("u_global")[:] = case.u0.ravel()
            self.oracles.append(o)
            self.cases.append(case)
        self.cells_per_step = sum(c.mesh.num_owned_cells for c in self.cases)
        self.dt = self.cases[0].dt

    def run(self, nsteps):
        """nsteps Euler steps on every core; returns wall seconds."""
        threads = [threading.Thread(target=o.step_euler, args=(self.dt, nsteps)) for o in self.oracles]
        t0 = time.perf_counter()
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        return time.perf_counter() - t0

    def sample_desc(self, nsteps):
        return (f"{self.cores} host threads x ({self.block_nx}x{self.block_nx}x2-triangle floodplain block, {nsteps} Euler steps), "
                f"oracle kind '{self.kind}' (reference src/swe/swe_petsc.c operator, serial per block, no MPI/PetscSF cost)")


def cpu_baseline(seconds, hr=False):
    arm = CpuArm(256, hr=hr)
    t1 = arm.run(2)  # warm-up + calibration
    nsteps = max(3, int(seconds / max(t1 / 2, 1e-6)))
    t = arm.run(nsteps)
    return {"value": arm.cells_per_step * nsteps / t, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample_desc(nsteps)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    arm = CpuArm(256, hr=args.hr)
    arm.run(max(1, min(args.warmup, 5)))
    t = arm.run(args.steps)
    value = arm.cells_per_step * args.steps / t
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "config C floodplain (bounded sample per step: %d cells on %d host threads)" % (arm.cells_per_step, arm.cores)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": arm.cores, "kind": arm.kind, "sample": arm.sample_desc(args.steps)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------------------
# GPU arm
# --------------------------------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist

    from rdycore_b200 import cases
    from rdycore_b200._lib import lib
    from rdycore_b200.operator import GATHER, TILED, TILED_ASYNC, TILED_FIXED, TILED_SPLIT, WARP, Operator, kernel_launch_count

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

    ny = args.ny or args.nx
    if args.tile is None:
        args.tile = int(os.environ.get("RDYB_TILE", "128" if args.second_order else "224"))
    px, py = block_grid(world)
    t_setup = time.perf_counter()
    if args.hr:
        case = cases.floodplain(nx=args.nx * px, ny=ny * py, eta0=12.0, hr=True, px=px, py=py, rank=rank, dam_column=(2000.0, 28.0))
    else:
        eta_default = 60.0 if args.second_order else 18.0
        case = cases.floodplain(nx=args.nx * px, ny=ny * py, eta0=eta_default if args.eta0 is None else args.eta0, px=px, py=py, rank=rank)
        if args.second_order:
            case.config.reconstruction = cases.MUSCL_MINMOD
    halo = case.mesh.halo() if world > 1 else None
    op = Operator.from_case(case, device=local_rank, halo=halo)
    kernel = {"auto": TILED_FIXED, "gather": GATHER, "tiled": TILED, "async": TILED_ASYNC, "fixed": TILED_FIXED, "warp": WARP, "split": TILED_SPLIT}[args.kernel]
    kernel_name = {GATHER: "gather", TILED: "tiled", TILED_ASYNC: "tiled_async", TILED_FIXED: "tiled_fixed", WARP: "warp", TILED_SPLIT: "tiled_split"}[kernel]
    op.set_kernel(kernel, args.tile)
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

    def barrier():
        op.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    # ---- device-resident timing: W warm-up steps, then exactly K steps between CUDA events on the operator's stream
    op.step_euler(case.dt, args.warmup)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    n0 = kernel_launch_count()
    op.step_euler(case.dt, args.steps)
    ms_per_step, _ = op.last_kernel_ms()  # event-synchronises
    kinfo = op.kernel_info()
    barrier()
    launches = kernel_launch_count() - n0
    clocks = sampler.stop()
    ms_per_step = max_over_ranks(ms_per_step)
    total_cells = sum_over_ranks(float(n_owned))
    value = total_cells / (ms_per_step * 1e-3)
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
        op.apply_host_ptr(case.dt, pu.value, pf.value)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.e2e_steps):
        op.apply_host_ptr(case.dt, pu.value, pf.value)  # returns after the D2H copy of F completed
    barrier()
    e2e_s = max_over_ranks((time.perf_counter() - t0) / args.e2e_steps)
    e2e = {"value": total_cells / e2e_s, "unit": UNIT, "h2d_bytes_per_step": int(sum_over_ranks(24.0 * nl)),
           "d2h_bytes_per_step": int(sum_over_ranks(24.0 * n_owned)), "ms_per_step": 1e3 * e2e_s,
           "call": "rdyb_operator_apply_host (PetscOperator apply shape: host u_local AoS in, host F AoS out, pinned)"}
    f_finite = bool(np.isfinite(f_host).all())
    lib().rdyb_host_free(pu)
    lib().rdyb_host_free(pf)

    if rank == 0:
        peak, peak_src = measured_peak()
        bytes_per_cell = 140 if args.hr else BYTES_PER_CELL_TRI
        if args.second_order:  # + gradients written and read (48 + 48) + per-slot LS weights (3 slots x 2 doubles) + face offsets (K1: 3 slots x 4; K2M: 1.5 edges x 4)
            bytes_per_cell += 96 + 48 + (96 if kernel == GATHER else 48)
        achieved = bytes_per_cell * n_owned / (ms_per_step * 1e-3) / 1e9  # this rank's kernel: per-GPU figure
        traffic = recorded_traffic(kernel_name + ("_hr" if args.hr else ""))
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {
                "workload": ("config D: HR floodplain" if args.hr else "config C: synthetic unstructured triangle floodplain")
                + f", {args.nx}x{ny}x2 = {2 * args.nx * ny} cells per GPU, bathymetry + Manning n + rain, reflecting walls, "
                + ("XQ2018" if False else "semi-implicit") + " friction, forward Euler"
                + (", second order (MUSCL, minmod), fully wet" if args.second_order else ""),
                "cells_total": int(total_cells), "decomposition": (f"{px}x{py} blocks, 1 ghost layer, per-step halo: " + ("border cells pushed into the neighbours' ghost slots over NVLink (CUDA IPC peer memory)" if args.halo == "p2p" else "NCCL send/recv")) if world > 1 else "single GPU",
                "kernel_info": kinfo,
                "kernel": kernel_name, "tile_cells": args.tile if kernel != GATHER else None, "courant_diagnostic": not args.no_courant,
                "l2": "inputs larger than L2 (state 240 MB + mesh > 1 GB per GPU vs 126 MB L2); no flush needed",
                "dt": case.dt, "setup_s": round(t_setup, 2),
            },
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": (traffic * n_owned if traffic else None), "peak_source": peak_src,
                         "algorithmic_bytes_per_cell": bytes_per_cell, "kernel": ("muscl_gradient_kernel + " + ("swe_gather_kernel<RECON>" if kernel == GATHER else "swe_tiled_muscl_kernel")) if args.second_order else "swe_%s_kernel" % kernel_name,
                         "note": "per-GPU: algorithmic bytes of one launch / CUDA-event time per launch on the launching stream"},
            "e2e": e2e,
            "gpu_launches": int(launches),
            "clocks": clocks,
            "checks": {"state_finite": finite, "f_finite": f_finite, "max_courant": courant},
        }
        if not args.no_cpu_baseline and world == 1:
            line["cpu_baseline"] = cpu_baseline(args.cpu_seconds, hr=args.hr)
        print(json.dumps(line), flush=True)
    op.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()
