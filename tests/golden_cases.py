"""Small seeded SWE cases shared by tests/golden/make_golden.py (which records the REFERENCE's results for them) and by
the parity tests (which replay them on the CPU oracles and on the CUDA path).

Each entry: name -> (builder() -> Case, stepper, nsteps).  Sizes are a few hundred cells so that the committed fixture
stays small; the physics options cover every branch of the path: both source methods, the
three boundary-condition types, hydrostatic reconstruction, wet/dry fronts, mixed quads + triangles, RK4.
"""
from __future__ import annotations

import numpy as np

from rdycore_b200 import cases
from rdycore_b200.mesh import HostMesh


def dam_break(source_method):
    def build():
        return cases.dam_break_quads(20, 20, source_method=source_method)
    return build


def dam_break_dry_bed():
    case = cases.dam_break_quads(30, 10, source_method=cases.SEMI_IMPLICIT)
    xc = case.mesh.array("cell_xc")
    case.u0[:, 0] = np.where(xc < 5.0, 2.0, 0.0)
    return case


def floodplain_rain():
    return cases.floodplain(nx=16, ny=12, eta0=8.0)


def floodplain_hr(source_method):
    def build():
        case = cases.floodplain(nx=16, ny=16, eta0=6.0, hr=True, source_method=source_method, dam_column=(60.0, 14.0))
        case.sources[:] = 0.0
        return case
    return build


def dirichlet_channel():
    mesh = HostMesh.quads(20, 10, 20.0, 10.0)
    bnd = cases.boundaries_from_predicates(mesh, [
        (1, "west", cases.DIRICHLET, lambda x, y: x < 1e-9),
        (2, "east", cases.CRITICAL_OUTFLOW, lambda x, y: x > 20.0 - 1e-9),
    ])
    n = mesh.num_owned_cells
    u0 = np.zeros((n, 3))
    u0[:, 0] = 1.0
    ne = len(bnd[0].edge_ids)
    dvals = np.tile(np.array([1.5, 0.75, 0.0]), (ne, 1)) * (1 + 0.1 * np.arange(ne)[:, None] / ne)
    return cases.Case("dirichlet_channel", mesh, cases.make_config(), bnd, u0, np.full(n, 0.03), None, 2e-3, dirichlet={0: dvals})


def mixed_elements():
    nx, ny = 12, 8
    xs, ys = np.meshgrid(np.linspace(0, 12.0, nx + 1), np.linspace(0, 8.0, ny + 1), indexing="xy")
    rng = np.random.default_rng(3)
    z = 0.3 * np.sin(xs) * np.cos(ys) + 0.05 * rng.random(xs.shape)
    xyz = np.stack([xs.ravel(), ys.ravel(), z.ravel()], axis=1)
    offs, verts = [0], []
    for j in range(ny):
        for i in range(nx):
            v0 = j * (nx + 1) + i
            quad = [v0, v0 + 1, v0 + nx + 2, v0 + nx + 1]
            if i < nx // 2:
                verts += quad
                offs.append(offs[-1] + 4)
            else:
                verts += [quad[0], quad[1], quad[2]]
                offs.append(offs[-1] + 3)
                verts += [quad[0], quad[2], quad[3]]
                offs.append(offs[-1] + 3)
    mesh = HostMesh.from_cells(xyz, np.array(offs, np.int32), np.array(verts, np.int32))
    bnd = cases.boundaries_from_predicates(mesh, [(1, "south", cases.CRITICAL_OUTFLOW, lambda x, y: y < 1e-9)])
    n = mesh.num_owned_cells
    xc = mesh.array("cell_xc")
    u0 = np.zeros((n, 3))
    u0[:, 0] = np.where(xc < 6.0, 2.0, 0.5)
    return cases.Case("mixed", mesh, cases.make_config(source_method=cases.IMPLICIT_XQ2018), bnd, u0, np.full(n, 0.02), None, 2e-3)


def moving_rhs_case():
    """A single RHS evaluation on a state with non-zero momenta and 3-component sources (flux divergence, primitive
    variables and the Courant diagnostic are recorded too)."""
    case = cases.floodplain(nx=14, ny=10, eta0=9.0)
    rng = np.random.default_rng(7)
    n = case.u0.shape[0]
    case.u0[:, 1] = rng.normal(0, 0.3, n) * case.u0[:, 0]
    case.u0[:, 2] = rng.normal(0, 0.3, n) * case.u0[:, 0]
    case.sources[:, 1] = 1e-4 * rng.normal(size=n)
    case.sources[:, 2] = 1e-4 * rng.normal(size=n)
    return case


def second_order(reconstruction, triangles):
    """numerics.second_order (MUSCL) on fully wet cases: the quad dam break, or the triangle floodplain under 40 m of water with a
    rough momentum field (the scheme has no dry-cell treatment, the reference returns NaN there)."""
    def build():
        if triangles:
            case = cases.floodplain(nx=20, ny=16, eta0=40.0)
            rng = np.random.default_rng(11)
            n = case.u0.shape[0]
            case.u0[:, 1] = rng.normal(0, 0.3, n) * case.u0[:, 0]
            case.u0[:, 2] = rng.normal(0, 0.3, n) * case.u0[:, 0]
        else:
            case = cases.dam_break_quads(nx=24, ny=12, source_method=cases.SEMI_IMPLICIT)
        case.config.reconstruction = reconstruction
        return case
    return build


def reference_deck(name):
    """the reference's own decks on its own mesh / IC / Manning / elevation files (tests/test_reference_decks.py)"""
    def build():
        from test_reference_decks import deck_case
        return deck_case(name)
    return build


# name -> (builder, "euler" | "rk4" | "rhs", nsteps)
GOLDEN = {
    "dam_break_xq2018": (dam_break(cases.IMPLICIT_XQ2018), "euler", 200),
    "dam_break_semi_implicit": (dam_break(cases.SEMI_IMPLICIT), "euler", 200),
    "dam_break_dry_bed": (dam_break_dry_bed, "euler", 300),
    "floodplain_rain": (floodplain_rain, "euler", 100),
    "floodplain_hr_semi_implicit": (floodplain_hr(cases.SEMI_IMPLICIT), "euler", 100),
    "floodplain_hr_xq2018": (floodplain_hr(cases.IMPLICIT_XQ2018), "euler", 100),
    "dirichlet_channel": (dirichlet_channel, "euler", 200),
    "mixed_elements": (mixed_elements, "euler", 200),
    "floodplain_rk4": (floodplain_rain, "rk4", 30),
    "moving_rhs": (moving_rhs_case, "rhs", 1),
    # second order: minmod and van Leer limiters, unlimited; Euler and RK4
    "muscl_minmod_quads": (second_order(cases.MUSCL_MINMOD, False), "euler", 150),
    "muscl_none_quads": (second_order(cases.MUSCL_NONE, False), "euler", 150),
    "muscl_vanleer_quads": (second_order(cases.MUSCL_VANLEER, False), "euler", 150),
    "muscl_minmod_triangles": (second_order(cases.MUSCL_MINMOD, True), "euler", 100),
    "muscl_vanleer_triangles_rk4": (second_order(cases.MUSCL_VANLEER, True), "rk4", 30),
    "muscl_minmod_rhs": (second_order(cases.MUSCL_MINMOD, True), "rhs", 1),
    # the reference's decks and fixture files
    "deck_ex2b_ic_file": (reference_deck("ex2b_ic_file"), "euler", 300),
    "deck_mixed_elements_ic_file": (reference_deck("mixed_elements_ic_file"), "euler", 300),
    "deck_four_mounds_60x24": (reference_deck("four_mounds_60x24"), "euler", 25),
    "deck_levee_hr": (reference_deck("levee.hr"), "euler", 200),
    "deck_houston1km": (reference_deck("Houston1km.DirichletBC"), "euler", 140),
}


def run_oracle(oracle, case, mode, nsteps):
    """Advance `oracle` (already loaded with `case`) and return the arrays the fixture records."""
    out = {}
    if mode == "euler":
        oracle.step_euler(case.dt, nsteps)
        out["u"] = oracle.array("u_global").copy()
    elif mode == "rk4":
        oracle.step_rk4(2 * case.dt, nsteps)
        out["u"] = oracle.array("u_global").copy()
    else:
        out["rhs"] = oracle.rhs(case.dt).copy()
        out["flux_divergence"] = oracle.array("flux_divergence").copy()
        out["primitive_variables"] = oracle.array("primitive_variables").copy()
        out["courant"] = np.array([oracle.courant()])
    for b, bd in enumerate(case.boundaries):
        if mode == "euler" and len(bd.edge_ids):
            out[f"bacc{b}"] = oracle.array("boundary_fluxes_accum", b).copy()
    return out
