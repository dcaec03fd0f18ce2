"""GPU: the second-order (MUSCL) path — least-squares gradients, limited face reconstruction, Roe on the reconstructed states
(ApplyInteriorFlux2R src/swe/swe_petsc.c:98-213; helpers src/operator_fluxes_ceed.c:884-1195) — against the CPU oracle.
Cases are fully wet: with dry cells the reference's second-order scheme itself drives depths negative and returns NaN
(pow(h, -1/3) in the friction term), so there is nothing to compare."""
import os

import numpy as np
import pytest

from rdycore_b200 import cases
from rdycore_b200.operator import GATHER, TILED, TILED_FIXED, Operator

from util import assert_state_close, make_oracle

pytestmark = pytest.mark.gpu


def _tiled_muscl(info: str) -> bool:
    """the second-order path ran on the tiled layout: K3M (bulk-async staged, the default) or K2M (RDYB_MUSCL_K2M=1 / shapes K3M cannot stage)"""
    return "K4M" in info or "K3M" in info or "K2M" in info
LIMITERS = [cases.MUSCL_MINMOD, cases.MUSCL_NONE, cases.MUSCL_VANLEER]


def _wet_floodplain(reconstruction, source_method=cases.SEMI_IMPLICIT):
    case = cases.floodplain(nx=40, ny=36, eta0=40.0, source_method=source_method)
    case.config.reconstruction = reconstruction
    rng = np.random.default_rng(1)
    case.u0[:, 1] = 0.3 * rng.normal(size=len(case.u0)) * case.u0[:, 0]
    case.u0[:, 2] = 0.3 * rng.normal(size=len(case.u0)) * case.u0[:, 0]
    return case


@pytest.mark.parametrize("reconstruction", LIMITERS)
@pytest.mark.parametrize("source_method", [cases.SEMI_IMPLICIT, cases.IMPLICIT_XQ2018])
def test_dam_break_quads_second_order(reconstruction, source_method):
    case = cases.dam_break_quads(nx=60, ny=60, source_method=source_method)
    case.config.reconstruction = reconstruction
    oracle = make_oracle(case)
    op = Operator.from_case(case)
    op.set_kernel(TILED_FIXED, 224)  # any tiled kind runs the second-order path on the tiled layout (K2M, swe_tiled_muscl.cuh)
    # Van Leer's smooth limiter amplifies round-off faster than minmod: the CPU port compiled with and without FMA contraction
    # (scratch/fma/sens.py, same C source) differs by 4e-12 / 9e-12 / 1.7e-11 after 100 / 200 / 400 steps on this case, the
    # GPU by 1.15e-11 after 400 — so the 1e-11 tolerance is checked at 150 steps there, 400 for minmod and unlimited.
    steps = 150 if reconstruction == cases.MUSCL_VANLEER else 400
    op.step_euler(case.dt, steps)
    oracle.step_euler(case.dt, steps)
    assert _tiled_muscl(op.kernel_info()) and "MUSCL)" in op.kernel_info()
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second-order dam break")
    first = cases.dam_break_quads(nx=60, ny=60, source_method=source_method)
    o1 = make_oracle(first)
    o1.step_euler(case.dt, steps)
    assert np.abs(o1.array("u_global") - oracle.array("u_global")).max() > 1e-2  # it really is a different scheme


@pytest.mark.parametrize("kernel", [GATHER, TILED])
@pytest.mark.parametrize("reconstruction", [cases.MUSCL_MINMOD, cases.MUSCL_VANLEER])
def test_wet_triangle_floodplain_second_order(reconstruction, kernel):
    case = _wet_floodplain(reconstruction)
    oracle = make_oracle(case)
    op = Operator.from_case(case)
    op.set_kernel(kernel, 128)
    if reconstruction == cases.MUSCL_VANLEER:
        # van Leer + forward Euler is unstable on this rough state (the oracle itself reaches NaN by step 200): RK4 there
        op.step_rk4(case.dt, 100)
        oracle.step_rk4(case.dt, 100)
    else:
        op.step_euler(case.dt, 200)
        oracle.step_euler(case.dt, 200)
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second-order floodplain")


@pytest.mark.parametrize("kernel", [GATHER, TILED])
def test_second_order_rhs_courant_and_rk4(kernel):
    case = _wet_floodplain(cases.MUSCL_MINMOD)
    oracle = make_oracle(case)
    op = Operator.from_case(case)
    op.set_kernel(kernel, 224)
    op.set_options(courant=True, flux_divergence=True)
    F_ref = oracle.rhs(case.dt).reshape(-1, 3).copy()
    F = op.apply_host(case.dt, oracle.array("u_local"))
    scale = np.abs(F_ref).max(axis=0)
    assert (np.abs(F - F_ref) / scale).max() < 1e-12
    fd = oracle.array("flux_divergence").reshape(-1, 3)
    assert (np.abs(op.get_field(0) - fd) / np.abs(fd).max(axis=0)).max() < 1e-12
    assert op.courant().max_courant_num == pytest.approx(oracle.courant(), rel=1e-12)
    assert _tiled_muscl(op.kernel_info()) == (kernel == TILED)
    d = op.courant()  # the arg of the maximum names an internal edge and one of its two cells, whichever kernel reported it
    ecells = case.mesh.view_arrays()["edge_cell_ids"].reshape(-1, 2)
    assert d.global_cell_id in ecells[d.global_edge_id] and ecells[d.global_edge_id].min() >= 0
    op.set_state(case.u0)
    op.step_rk4(case.dt, 40)
    oracle.array("u_global")[:] = case.u0.ravel()
    oracle.step_rk4(case.dt, 40)
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second-order RK4")
    # bitwise repeatable
    op2 = Operator.from_case(case)
    op2.set_kernel(kernel, 224)
    op2.step_rk4(case.dt, 40)
    assert np.array_equal(op2.get_state(), op.get_state())


@pytest.mark.parametrize("reconstruction", LIMITERS)
def test_tiled_and_gather_second_order_agree(reconstruction, monkeypatch):
    """K2M solves every edge once from shared memory, K1 twice from global memory: same face states, same fluxes; only the
    association of flux * length / area differs, so one right-hand side agrees to round-off.  RDYB_MUSCL_K1 keeps K1 on a tiled layout."""
    case = _wet_floodplain(reconstruction)
    ops = []
    for kind, env in ((TILED, None), (GATHER, None), (TILED_FIXED, "1")):
        if env:
            monkeypatch.setenv("RDYB_MUSCL_K1", env)
        op = Operator.from_case(case)
        op.set_kernel(kind, 160)
        op.set_options(courant=True)
        ops.append(op)
    oracle = make_oracle(case)
    oracle.rhs(case.dt)  # fills u_local from the initial state
    u_local = oracle.array("u_local")
    assert np.abs(u_local.reshape(-1, 3)[:, 1:]).max() > 0.0
    F = [op.apply_host(case.dt, u_local) for op in ops]
    assert _tiled_muscl(ops[0].kernel_info()) and "K1, MUSCL" in ops[1].kernel_info() and "K1, MUSCL" in ops[2].kernel_info()
    scale = np.abs(F[1]).max(axis=0)
    assert (np.abs(F[0] - F[1]) / scale).max() < 1e-14
    assert np.array_equal(F[1], F[2])
    c = [op.courant().max_courant_num for op in ops]
    assert c[0] == pytest.approx(c[1], rel=1e-14) and c[1] == c[2]


def test_hr_with_second_order_is_refused_like_the_reference():
    """CreateOperator: "-second_order cannot be used with well_balancing = HR simultaneously" (src/operator.c:388-389)."""
    case = cases.floodplain(nx=32, ny=28, eta0=9.0, hr=True)
    case.config.reconstruction = cases.MUSCL_MINMOD
    with pytest.raises(RuntimeError, match="second_order cannot be used with well_balancing = HR"):
        Operator.from_case(case)
    with pytest.raises(RuntimeError):
        make_oracle(case)


@pytest.mark.parametrize("reconstruction", LIMITERS)
@pytest.mark.parametrize("tile", [128, 224])
def test_async_tiled_second_order(reconstruction, tile, monkeypatch):
    """K3M = K2M's arithmetic behind K3's bulk-async staging: the same right-hand side as K2M to round-off (same face states, same
    flux * length, same per-cell order), the oracle's state after 100 steps, the same Courant number."""
    case = _wet_floodplain(reconstruction)
    oracle = make_oracle(case)
    oracle.rhs(case.dt)
    u_local = oracle.array("u_local").copy()
    monkeypatch.setenv("RDYB_MUSCL_K2M", "1")
    ref = Operator.from_case(case)
    ref.set_kernel(TILED, tile)
    ref.set_options(courant=True)
    F_ref = ref.apply_host(case.dt, u_local)
    monkeypatch.delenv("RDYB_MUSCL_K2M")
    op = Operator.from_case(case)
    op.set_kernel(TILED, tile)
    op.set_options(courant=True)
    F = op.apply_host(case.dt, u_local)
    assert "K3M" in op.kernel_info() and "K2M" in ref.kernel_info()
    assert (np.abs(F - F_ref) / np.abs(F_ref).max(axis=0)).max() < 1e-14  # the same expressions, compiled in two kernels
    assert op.courant().max_courant_num == pytest.approx(ref.courant().max_courant_num, rel=1e-14)
    # unlimited extrapolation + forward Euler is unstable in the reference scheme itself (the oracle reaches NaN by step 100)
    steps = 30 if reconstruction == cases.MUSCL_NONE else 100
    if reconstruction == cases.MUSCL_VANLEER:
        op.step_rk4(case.dt, steps); oracle.step_rk4(case.dt, steps)
    else:
        op.step_euler(case.dt, steps); oracle.step_euler(case.dt, steps)
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second-order floodplain on K3M")


@pytest.mark.parametrize("cap", [0, 3], ids=["fullgrid", "cap3"])
@pytest.mark.parametrize("consumers", [192, 256])
@pytest.mark.parametrize("reconstruction", LIMITERS)
def test_fixed_shape_second_order_k4m(reconstruction, consumers, cap, monkeypatch):
    """K4M on 192-cell tiles (swe_tiled_muscl_fixed.cuh).  Mode 2 computes the least-squares gradients of tile AND halo cells inside the
    tile kernel from a two-ring halo with the instruction sequence of muscl_gradient_kernel; mode 0 (RDYB_K4M_GRAD_ARRAY=1) reads the
    gradient array that kernel writes.  The same gradients bit for bit => the same right-hand side bit for bit.  Against K3M
    (RDYB_MUSCL_K3M=1) the right-hand side agrees to round-off (K4M takes u, v, sqrt(h) of a face from one MUFU seed); against the oracle
    the state after 100 steps, with the full grid and with 3 CTAs walking 5 tiles each through both stages."""
    case = _wet_floodplain(reconstruction)
    oracle = make_oracle(case)
    oracle.rhs(case.dt)
    u_local = oracle.array("u_local").copy()
    monkeypatch.setenv("RDYB_K4M_CONSUMERS", str(consumers))

    def make(env=None):
        if env:
            monkeypatch.setenv(env, "1")
        op = Operator.from_case(case)
        op.set_kernel(TILED, 192)
        op.set_options(courant=True)
        if cap:
            op.set_max_ctas(cap)
        F = op.apply_host(case.dt, u_local)
        if env:
            monkeypatch.delenv(env)
        return op, F

    ref, F_ref = make("RDYB_MUSCL_K3M")
    arr, F_arr = make("RDYB_K4M_GRAD_ARRAY")
    op, F = make()
    assert "K3M" in ref.kernel_info() and "K4M" in arr.kernel_info() and "muscl_gradient_kernel" in arr.kernel_info()
    assert "K4M" in op.kernel_info() and "in-kernel gradients" in op.kernel_info()
    assert np.array_equal(F, F_arr)  # in-kernel gradients == the gradient kernel's, bit for bit
    assert op.courant().max_courant_num == arr.courant().max_courant_num
    assert (np.abs(F - F_ref) / np.abs(F_ref).max(axis=0)).max() < 1e-13
    assert op.courant().max_courant_num == pytest.approx(ref.courant().max_courant_num, rel=1e-13)
    steps = 30 if reconstruction == cases.MUSCL_NONE else 100  # see test_async_tiled_second_order
    if reconstruction == cases.MUSCL_VANLEER:
        op.step_rk4(case.dt, steps); oracle.step_rk4(case.dt, steps)
    else:
        op.step_euler(case.dt, steps); oracle.step_euler(case.dt, steps)
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second-order floodplain on K4M")
    for o in (ref, arr, op):
        o.close()


def test_k4m_conserves_mass_and_repeats_bitwise():
    """Both tiles of a shared edge reconstruct from bit-identical gradients (each computed in its own stage): with closed walls and no
    rain the mass is conserved to round-off, and the result does not depend on which CTA ran which tile."""
    case = _wet_floodplain(cases.MUSCL_MINMOD)
    case.sources[:] = 0.0
    runs = []
    for cap in (0, 3, 0):
        op = Operator.from_case(case)
        op.set_kernel(TILED, 192)
        if cap:
            op.set_max_ctas(cap)
        m0 = op.mass()
        op.step_euler(case.dt, 100)
        m1 = op.mass()
        assert "in-kernel gradients" in op.kernel_info()
        assert abs(m1 - m0) <= 1e-12 * m0 * np.sqrt(100)
        runs.append(op.get_state())
        op.close()
    assert np.array_equal(runs[0], runs[1]) and np.array_equal(runs[0], runs[2])


@pytest.mark.parametrize("reconstruction", [cases.MUSCL_MINMOD, cases.MUSCL_VANLEER])
def test_k4m_on_a_small_quad_mesh_uses_the_gradient_array(reconstruction):
    """Quads have eight least-squares weights per cell, K4M stages six: a quad mesh small enough for K4M's edge capacity (a single
    192-cell tile holds at most 352 edges) runs it with the gradient array - four cell_edges planes, no halo at all."""
    case = cases.dam_break_quads(nx=12, ny=12, source_method=cases.SEMI_IMPLICIT)
    case.config.reconstruction = reconstruction
    oracle = make_oracle(case)
    op = Operator.from_case(case)
    op.set_kernel(TILED, 192)
    op.step_euler(case.dt, 60)
    oracle.step_euler(case.dt, 60)
    info = op.kernel_info()
    assert "K4M" in info and "muscl_gradient_kernel" in info
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second-order quads on K4M")
    op.close()


def test_k4m_single_tile_without_halo():
    """A triangle mesh that fits one tile: no first ring, no second ring (zero-byte index slices), in-kernel gradients all the same."""
    case = cases.floodplain(nx=9, ny=8, eta0=40.0)
    case.config.reconstruction = cases.MUSCL_MINMOD
    oracle = make_oracle(case)
    op = Operator.from_case(case)
    op.set_kernel(TILED, 192)
    op.step_euler(case.dt, 60)
    oracle.step_euler(case.dt, 60)
    assert "in-kernel gradients" in op.kernel_info()
    assert_state_close(op.get_state(), oracle.array("u_global"), what="second order, one tile")
    op.close()
