"""Shared helpers of the parity tests: build an oracle from a Case, compare states with the stated fp64 tolerance."""
from __future__ import annotations

import numpy as np

from oracle import binding

REL_TOL = 1e-11  # BASELINE.json north_star: max relative error 1e-11 in h, hu, hv
ABS_TOL_DRY = 1e-12


def make_oracle(case, kind=None):
    kind = kind or binding.best_kind()
    arr, nb = case.boundary_views()
    view = case.mesh.view()
    o = binding.Oracle(kind, case.config, view, arr if nb else None)
    o._case_keep = (case, arr, view)
    o.array("material_properties")[:] = case.mannings
    if case.sources is not None:
        o.array("external_sources")[:] = case.sources.ravel()
    for b, vals in case.dirichlet.items():
        o.array("boundary_values", b)[:] = np.asarray(vals).ravel()
    o.array("u_global")[:] = case.u0.ravel()
    return o


def state_error(u_gpu: np.ndarray, u_ref: np.ndarray, tiny_h: float = 1e-7):
    """SURVEY §8c tolerance: max over owned cells of |x_gpu - x_ref| / max(|x_ref|, 1e-6 max|x|) per component for cells with
    h_ref >= 10 tiny_h; cells below that are compared absolutely.  Returns (max_rel, max_abs_dry, n_wetdry_mismatch)."""
    u_gpu = np.asarray(u_gpu).reshape(-1, 3)
    u_ref = np.asarray(u_ref).reshape(-1, 3)
    wet = u_ref[:, 0] >= 10 * tiny_h
    max_rel = 0.0
    for c in range(3):
        floor = 1e-6 * max(np.abs(u_ref[:, c]).max(), 1e-300)
        denom = np.maximum(np.abs(u_ref[wet, c]), floor)
        if wet.any():
            max_rel = max(max_rel, float((np.abs(u_gpu[wet, c] - u_ref[wet, c]) / denom).max()))
    max_abs = float(np.abs(u_gpu[~wet] - u_ref[~wet]).max()) if (~wet).any() else 0.0
    mismatch = int(((u_gpu[:, 0] < tiny_h) != (u_ref[:, 0] < tiny_h)).sum())
    return max_rel, max_abs, mismatch


def assert_state_close(u_gpu, u_ref, tiny_h=1e-7, rel=REL_TOL, abs_dry=ABS_TOL_DRY, what=""):
    assert not np.isnan(np.asarray(u_ref)).any(), f"{what}: oracle produced NaN (bad test case)"
    assert not np.isnan(np.asarray(u_gpu)).any(), f"{what}: GPU produced NaN"
    max_rel, max_abs, mismatch = state_error(u_gpu, u_ref, tiny_h)
    assert max_rel <= rel, f"{what}: max relative error {max_rel:.3e} > {rel:.1e}"
    assert max_abs <= abs_dry, f"{what}: max absolute error on near-dry cells {max_abs:.3e} > {abs_dry:.1e}"
    return max_rel, max_abs, mismatch
