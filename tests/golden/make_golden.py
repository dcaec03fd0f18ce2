#!/usr/bin/env python
"""Record golden vectors of the SWE path from the REFERENCE's own code.

Run in the build container (where /root/reference exists):   python tests/golden/make_golden.py
It builds oracle/_ref/libswe_oracle_ref.so (the reference's src/swe/swe_petsc.c, swe_roe_flux_petsc.h, petsc_operator.c and
operator_{fluxes,sources}_petsc.c compiled where they lie, oracle/Makefile) and runs every case of tests/golden_cases.py
through it.  The results go to tests/golden/swe_golden.npz, which is committed: /root/reference does not exist on the GPU
box, the fixture does.  Each record also stores the case's inputs' checksum so that a drifting case definition is noticed.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def input_digest(case) -> str:
    h = hashlib.sha256()
    va = case.mesh.view_arrays()
    for k in ("cell_areas", "cell_dz_dx", "cell_dz_dy", "edge_cell_ids", "edge_lengths", "edge_cn", "edge_sn", "internal_edge_ids"):
        h.update(np.ascontiguousarray(va[k]).tobytes())
    h.update(np.ascontiguousarray(case.u0).tobytes())
    h.update(np.ascontiguousarray(case.mannings).tobytes())
    if case.sources is not None:
        h.update(np.ascontiguousarray(case.sources).tobytes())
    h.update(np.float64(case.dt).tobytes())
    return h.hexdigest()


def main():
    from golden_cases import GOLDEN, run_oracle
    from oracle import binding
    from util import make_oracle

    if not os.path.isdir("/root/reference/src/swe"):
        raise SystemExit("make_golden.py needs the reference tree at /root/reference")
    binding.build("reference")
    out = {}
    for name, (builder, mode, nsteps) in GOLDEN.items():
        case = builder()
        oracle = make_oracle(case, kind="reference")
        assert oracle.kind == "reference"
        rec = run_oracle(oracle, case, mode, nsteps)
        for k, v in rec.items():
            out[f"{name}/{k}"] = v
        out[f"{name}/digest"] = np.frombuffer(input_digest(case).encode(), dtype=np.uint8)
        u = rec.get("u", rec.get("rhs"))
        print(f"{name:32s} {mode:5s} x{nsteps:4d}  cells {case.u0.shape[0]:5d}  |x|max {np.nanmax(np.abs(u)):.6g}  nan {int(np.isnan(u).sum())}")
    path = os.path.join(HERE, "swe_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
