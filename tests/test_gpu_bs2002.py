"""Bradford-Sanders (2002) well balancing with the semantics of the reference's libCEED operator.

Oracle: oracle/_ref/libbs2002_oracle_ref.so - the reference's own Q-function headers (SWEEtaVertex, src/swe/swe_well_balancing.h:30;
SWEFlux_Roe + ComputeDhv + the boundary Q-functions, src/swe/swe_fluxes_ceed.h; SWERiemannFlux_Roe, swe_roe_flux_ceed.h; the source
Q-functions, swe_sources_ceed.h) compiled in place behind oracle/ceed_qf_shim and driven edge by edge (oracle/bs2002_ref_driver.c).
CUDA path: rdyb_operator_enable_bs2002 = bs2002_eta_vertex_kernel before every RHS + the g dhv^2 / 24 term in K1.

CPU part (no GPU): the oracle's plumbing is pinned against the PETSc-operator oracle with the term switched off (the two backends
then differ by round-off only on wet cells), and the term does what it is for: it changes the momentum balance of partially
submerged cells and vanishes where every vertex of an edge is equally deep.
"""
import os

import numpy as np
import pytest

from oracle import bs2002_binding
from rdycore_b200 import cases, io
from rdycore_b200.operator import Operator

from util import assert_state_close, make_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "ref_fixtures")
needs_oracle = pytest.mark.skipif(not (bs2002_binding.available() or os.path.isdir("/root/reference/src/swe")),
                                  reason="oracle/_ref/libbs2002_oracle_ref.so not built (needs /root/reference at build time)")


def _bowl(source_method=cases.IMPLICIT_XQ2018):
    """driver/tests/swe_roe/parabolic_bowl.yaml - the deck of the reference's swe_bs2002 CTest (driver/tests/swe_roe/CMakeLists.txt:238-250):
    6400 triangles over a parabolic bed, water released from rest in one half, Manning 0.033, reflecting walls."""
    ex = io.read_exodus(os.path.join(FIX, "parabolic_bowl.exo"))
    mesh = io.mesh_from_exodus(ex)
    n = mesh.num_owned_cells
    u0 = io.read_petsc_vec(os.path.join(FIX, "parabolic_bowl.ic.int32.bin")).reshape(-1, 3)
    assert u0.shape == (n, 3)
    bnd = io.boundaries_from_side_sets(mesh, ex, bc_types={})
    cfg = cases.make_config(source_method=source_method)
    return cases.Case("parabolic_bowl", mesh, cfg, bnd, u0, np.full(n, 0.033), np.zeros((n, 3)), 0.01)


def _shore(eta0=8.0, **kw):
    """floodplain triangles with a shoreline: many cells have one or two vertices above the water"""
    case = cases.floodplain(nx=40, ny=32, eta0=eta0, **kw)
    rng = np.random.default_rng(3)
    case.u0[:, 1] = 0.2 * rng.normal(size=case.u0.shape[0]) * case.u0[:, 0]
    case.u0[:, 2] = 0.2 * rng.normal(size=case.u0.shape[0]) * case.u0[:, 0]
    return case


@needs_oracle
def test_qfunction_oracle_without_the_term_agrees_with_the_petsc_operator_oracle():
    """With eta_vertices left at the vertex elevations (WELL_BALANCING_NONE, src/operator_fluxes_ceed.c:687-695) dhv is 0 and the
    libCEED operator is the PETSc operator up to round-off (sqrt vs pow, the order of the flux sum) on a fully wet state: this pins
    the gather / scatter plumbing of bs2002_ref_driver.c against the oracle that is pinned by the reference's own code."""
    case = _shore(eta0=40.0)
    ceed = bs2002_binding.Bs2002Oracle(case, bs2002=False)
    petsc = make_oracle(case)
    F_ceed = ceed.rhs(case.dt).copy()
    F_petsc = petsc.rhs(case.dt).reshape(-1, 3)
    assert (np.abs(F_ceed - F_petsc) / np.abs(F_petsc).max(axis=0)).max() < 1e-13
    assert ceed.courant() == pytest.approx(petsc.courant(), rel=1e-12)
    ceed.step_euler(case.dt, 40)
    petsc.step_euler(case.dt, 40)
    assert_state_close(ceed.u, petsc.array("u_global"), what="Q-function oracle vs PETSc oracle")


@needs_oracle
def test_the_term_acts_on_partially_submerged_cells_only():
    case = _shore(eta0=8.0)
    on, off = bs2002_binding.Bs2002Oracle(case, bs2002=True), bs2002_binding.Bs2002Oracle(case, bs2002=False)
    F_on, F_off = on.rhs(case.dt).copy(), off.rhs(case.dt).copy()
    changed = np.abs(F_on - F_off).max(axis=1) > 1e-12
    assert 50 < changed.sum() < 0.8 * len(changed)
    assert np.array_equal(F_on[:, 0], F_off[:, 0])  # the mass equation is untouched (src/swe/swe_roe_flux_ceed.h:119-121)
    # eta at the vertices: the lake level wherever the surrounding cells are fully submerged (a dry cell contributes its LOWEST
    # vertex elevation, so elsewhere eta may lie below the vertex itself)
    z = case.mesh.vertices()[:, 2]
    low = z < 3.0
    assert low.sum() > 100 and np.allclose(on.eta[low], 8.0, rtol=0, atol=1e-9)


@pytest.mark.gpu
@needs_oracle
@pytest.mark.parametrize("name", ["parabolic_bowl_xq2018", "parabolic_bowl_semi_implicit", "shoreline"])
def test_cuda_bs2002_against_the_reference_qfunctions(name):
    case = _shore(eta0=8.0) if name == "shoreline" else _bowl(cases.SEMI_IMPLICIT if "semi" in name else cases.IMPLICIT_XQ2018)
    oracle = bs2002_binding.Bs2002Oracle(case, bs2002=True)
    op = Operator.from_case(case)
    op.enable_bs2002()
    op.set_options(courant=True)
    # one right-hand side: F, the vertex elevations, the Courant number
    F_ref = oracle.rhs(case.dt).copy()
    F = op.apply(case.dt)
    scale = np.maximum(np.abs(F_ref).max(axis=0), 1e-300)
    assert (np.abs(F - F_ref) / scale).max() < 1e-12
    assert np.abs(op.eta_vertices() - oracle.eta).max() < 1e-13 * max(1.0, np.abs(oracle.eta).max())
    assert op.courant().max_courant_num == pytest.approx(oracle.courant(), rel=1e-12)
    # and it is not the operator without the term
    plain = Operator.from_case(case)
    assert np.abs(plain.apply(case.dt) - F_ref).max() > 1e-8
    # time stepping.  The oracle is the libCEED arithmetic (sqrt, strict wet tests, left-then-right flux sums), the kernel mirrors the
    # PETSc operator's; across the bowl's moving shoreline the two orders of operations drift apart faster than two builds
    # of one source do: 1e-11 after 100 steps, 1e-10 after 300 (measured on the B200: 1.6e-11 / 2.2e-11 after 300).  On the random
    # bathymetry of the floodplain the reference's scheme itself reaches NaN within 10 steps with the term on: one RHS only there.
    if name != "shoreline":
        for steps, tol in ((100, 1e-11), (200, 1e-10)):
            op.step_euler(case.dt, steps)
            oracle.step_euler(case.dt, steps)
            u = op.get_state()
            assert not np.isnan(oracle.u).any() and np.abs(oracle.u[:, 1:]).max() > 1e-4
            assert_state_close(u, oracle.u, rel=tol, what=f"BS2002 {name} after {steps} more steps")
    assert "K1" in op.kernel_info()


@pytest.mark.gpu
def test_bs2002_needs_triangles_and_a_serial_mesh():
    from rdycore_b200._lib import RdybError

    op = Operator.from_case(cases.dam_break_quads(10, 10))
    with pytest.raises(RdybError, match="triangles"):
        op.enable_bs2002()
