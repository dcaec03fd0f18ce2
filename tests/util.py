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


def celerity_momentum(u_ref) -> float:
    """h sqrt(g h) at the deepest cell: the momentum a unit-Froude flow would carry.  The scale of momentum round-off in states that
    are (nearly) at rest, where the momentum itself is ~0 but every step still differences pressure fluxes g h^2 / 2."""
    h = float(np.abs(np.asarray(u_ref).reshape(-1, 3)[:, 0]).max())
    return h * np.sqrt(9.806 * h)


def state_error(u_gpu: np.ndarray, u_ref: np.ndarray, tiny_h: float = 1e-7, qscale_floor: float = 0.0, rel_floor: float = 1e-3, scales=None):
    """Parity metric.  Per component x in {h, hu, hv}: max over owned cells with h_ref >= 10 tiny_h of
    |x_gpu - x_ref| / max(|x_ref|, floor), with floor = 1e-3 * (max h for h; max momentum magnitude sqrt(hu^2+hv^2) for hu
    and hv, which mix through the edge-normal rotation).  Cells below 10 tiny_h are compared absolutely.
    Why the floor is 1e-3 and not SURVEY §8c's 1e-6: a momentum component that crosses zero carries the absolute round-off
    of the dominant component (~1e-15 * scale), so its pointwise relative error at |x| = 1e-6 * scale is ~1e-9 for ANY
    implementation that is not bit-identical (measured on the B200: 3e-10 after 10 steps even with FMA contraction off, while
    the norm-wise error max|dx| / max|x| stays at 2e-15; see DESIGN.md "Parity").
    `scales` = (max h, max momentum magnitude) of the WHOLE mesh when u_ref is one rank's part of it (a rank whose block is a lake at
    rest would otherwise measure round-off against its own ~1e-9 momenta).
    `rel_floor` selects the floor (1e-3 here; pass 1e-6 to get SURVEY's figure - bench.py's checks.parity reports both).
    Returns (max_rel, max_abs_dry, n_wetdry_mismatch, normwise) with normwise = max_c max|dx_c| / scale_c."""
    u_gpu = np.asarray(u_gpu).reshape(-1, 3)
    u_ref = np.asarray(u_ref).reshape(-1, 3)
    wet = u_ref[:, 0] >= 10 * tiny_h
    hscale = max(float(np.abs(u_ref[:, 0]).max()), 1e-300)
    qscale = max(float(np.sqrt(u_ref[:, 1] ** 2 + u_ref[:, 2] ** 2).max()), qscale_floor, 1e-300)
    if scales is not None:
        hscale, qscale = max(float(scales[0]), 1e-300), max(float(scales[1]), qscale_floor, 1e-300)
    max_rel = normwise = 0.0
    for c, scale in ((0, hscale), (1, qscale), (2, qscale)):
        diff = np.abs(u_gpu[:, c] - u_ref[:, c])
        normwise = max(normwise, float(diff.max()) / scale)
        if wet.any():
            denom = np.maximum(np.abs(u_ref[wet, c]), rel_floor * scale)
            max_rel = max(max_rel, float((diff[wet] / denom).max()))
    max_abs = float(np.abs(u_gpu[~wet] - u_ref[~wet]).max()) if (~wet).any() else 0.0
    mismatch = int(((u_gpu[:, 0] < tiny_h) != (u_ref[:, 0] < tiny_h)).sum())
    return max_rel, max_abs, mismatch, normwise


def assert_state_close(u_gpu, u_ref, tiny_h=1e-7, rel=REL_TOL, abs_dry=ABS_TOL_DRY, what="", qscale_floor=0.0):
    assert not np.isnan(np.asarray(u_ref)).any(), f"{what}: oracle produced NaN (bad test case)"
    assert not np.isnan(np.asarray(u_gpu)).any(), f"{what}: GPU produced NaN"
    max_rel, max_abs, mismatch, normwise = state_error(u_gpu, u_ref, tiny_h, qscale_floor)
    assert max_rel <= rel, f"{what}: max relative error {max_rel:.3e} > {rel:.1e}"
    assert normwise <= 1e-13, f"{what}: norm-wise error {normwise:.3e} > 1e-13"
    assert max_abs <= abs_dry, f"{what}: max absolute error on near-dry cells {max_abs:.3e} > {abs_dry:.1e}"
    return max_rel, max_abs, mismatch
