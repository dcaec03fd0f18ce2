"""Secondary measurement: one forcing update per coupling interval on config C (10 M triangles, one B200).

  host path   = what RDyApplyForcing does: a host loop maps the raster onto the owned cells (timed with the reference's own
                RDyForcingSetRasterData through oracle/_ref, one thread), then RDySetRegionalWaterSource uploads num_owned doubles;
  device path = rdyb_operator_set_source_from_dataset: the raster goes up (KBs to MBs), one gather kernel writes the source array.
Also the one-off cost of the nearest-neighbour map (reference: every pixel against every cell on the host; here a window scan on the
GPU), the CPU figure measured on a sample of cells and scaled.  Prints one JSON line; run on the GPU box:
    python tools/bench_forcing.py [--nx 2236] [--raster 512]"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import forcing_binding  # noqa: E402  (the checker and the timed CPU baseline, as in bench.py's cpu_baseline leg)
from rdycore_b200 import cases, forcing  # noqa: E402
from rdycore_b200.operator import Operator  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=2236)
    ap.add_argument("--raster", type=int, default=512)
    ap.add_argument("--reps", type=int, default=10)
    a = ap.parse_args()
    case = cases.floodplain(nx=a.nx, ny=a.nx, eta0=18.0)
    n = case.mesh.num_owned_cells
    xc, yc = case.mesh.array("cell_xc")[:n].copy(), case.mesh.array("cell_yc")[:n].copy()
    op = Operator.from_case(case)
    oracle = forcing_binding.ForcingOracle()
    R = a.raster
    cs = max(np.ptp(xc), np.ptp(yc)) / (R - 2)
    xlc, ylc = xc.min() - cs, yc.min() - cs
    rng = np.random.default_rng(1)
    vec = np.concatenate([[R, R, xlc, ylc, cs], rng.uniform(0.0, 30.0, R * R)])
    # --- the map ---
    t0 = time.perf_counter()
    dmap = forcing.DatasetMap.raster(op, xc, yc, R, R, xlc, ylc, cs)
    op.synchronize()
    t_map_gpu = time.perf_counter() - t0
    sample = np.linspace(0, n - 1, 64).astype(int)
    t0 = time.perf_counter()
    idx_s = oracle.map_raster(xc[sample], yc[sample], R, R, xlc, ylc, cs)
    t_map_cpu_sample = time.perf_counter() - t0
    idx = dmap.indices()
    assert np.array_equal(idx[sample], idx_s)
    # --- one update ---
    t_host_loop, t_host_upload, t_dev = [], [], []
    for _ in range(a.reps):
        t0 = time.perf_counter(); rain = oracle.set_raster(vec, 5, idx); t1 = time.perf_counter()
        op.set_external_source(0, rain); op.synchronize(); t2 = time.perf_counter()
        forcing.set_source_from_dataset(op, 0, dmap, vec, offset=5, stride=1, scale=forcing.MM_PER_HR_TO_M_PER_S); op.synchronize(); t3 = time.perf_counter()
        t_host_loop.append(t1 - t0); t_host_upload.append(t2 - t1); t_dev.append(t3 - t2)
    assert np.array_equal(op.get_external_source(0), rain)
    med = lambda v: float(np.median(v))  # noqa: E731
    print(json.dumps({
        "what": "forcing update of the water source from a rain raster, per coupling interval", "cells": int(n), "raster": [R, R],
        "host_path_ms": {"map_loop_reference_code_1_thread": 1e3 * med(t_host_loop), "upload_and_permute": 1e3 * med(t_host_upload),
                         "total": 1e3 * (med(t_host_loop) + med(t_host_upload))},
        "device_path_ms": 1e3 * med(t_dev), "speedup": (med(t_host_loop) + med(t_host_upload)) / med(t_dev),
        "h2d_bytes": {"host_path": 8 * int(n), "device_path": 8 * int(vec.size)},
        "map_build_s": {"device_all_cells": t_map_gpu, "reference_code_cpu_scaled_from_64_cells": t_map_cpu_sample / 64 * n},
    }))


if __name__ == "__main__":
    main()
