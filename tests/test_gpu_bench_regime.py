"""GPU parity in the regime the bench runs in: many tiles per persistent CTA, meshes of >= 1 M cells.

The other parity tests use meshes of <= 25 600 cells, i.e. <= 115 tiles against a persistent grid of 444 CTAs: ONE tile per CTA.
bench.py runs ~100 tiles per CTA (10 M cells).  Here a 1.04 M-cell floodplain (4 629 tiles of 224 cells, >= 10 per CTA on a
B200, and 1 543 per CTA with the grid capped at 3) is stepped 20 times and compared with the CPU oracle (the reference's own
swe_petsc.c compiled in place, else the bit-identical port) at the tolerance of tests/util.py; plus what DESIGN.md section 5
claims at full size: bitwise repeatability and mass conservation to round-off.
"""
import numpy as np
import pytest

from rdycore_b200 import cases
from rdycore_b200.operator import TILED, TILED_FIXED, Operator

from util import assert_state_close, make_oracle, state_error

pytestmark = pytest.mark.gpu

NX = 720  # 720 x 720 x 2 = 1 036 800 triangles
STEPS = 20


def _case(hr=False, second_order=False):
    if second_order:
        case = cases.floodplain(nx=NX, ny=NX, eta0=40.0)  # fully wet: the second-order scheme has no dry-cell treatment
        case.config.reconstruction = cases.MUSCL_MINMOD
        g = case.mesh.array("cell_global_ids")[case.mesh.array("owned_to_local")].astype(np.float64)
        case.u0[:, 1] = 0.3 * np.sin(0.37 * g) * case.u0[:, 0]
        case.u0[:, 2] = 0.3 * np.cos(0.23 * g) * case.u0[:, 0]
        return case
    if hr:
        return cases.floodplain(nx=NX, ny=NX, eta0=9.0, hr=True, dam_column=(600.0, 24.0))
    return cases.floodplain(nx=NX, ny=NX, eta0=12.0)


_ORACLE_CACHE = {}


def _reference_state(key, case):
    """the oracle's state after STEPS steps (8 s of CPU per case: computed once per variant)"""
    if key not in _ORACLE_CACHE:
        oracle = make_oracle(case)
        oracle.step_euler(case.dt, STEPS)
        _ORACLE_CACHE[key] = (oracle.array("u_global").reshape(-1, 3).copy(), oracle.courant())
        oracle.close()
    return _ORACLE_CACHE[key]


@pytest.mark.parametrize("cap", [0, 3, 148], ids=["fullgrid", "cap3", "cap148"])
@pytest.mark.parametrize("variant", ["first_order", "hr"])
def test_k4_one_million_cells_against_the_oracle(variant, cap):
    case = _case(hr=variant == "hr")
    assert case.mesh.num_owned_cells >= 1_000_000
    u_ref, c_ref = _reference_state(variant, case)
    op = Operator.from_case(case)
    op.set_kernel(TILED_FIXED, 224)
    if cap:
        op.set_max_ctas(cap)
    op.step_euler(case.dt, STEPS - 1)
    op.reset_diagnostics()  # the oracle's Courant number is the one of its last RHS evaluation
    op.step_euler(case.dt, 1)
    u = op.get_state()
    assert "K4" in op.kernel_info()
    rel, _, mismatch = assert_state_close(u, u_ref, what=f"1 M cells {variant} cap={cap}")
    assert mismatch == 0
    assert op.courant().max_courant_num == pytest.approx(c_ref, rel=1e-12)
    wet = (case.u0[:, 0] > 0.0).mean()
    assert 0.2 < wet < 0.95  # the shoreline (and, for HR, the dam-break front) lies inside the mesh: wet/dry edges are exercised
    assert np.abs(u_ref[:, 1]).max() > 1e-3  # and the water moves
    op.close()


@pytest.mark.parametrize("tile", [128, 192], ids=["K3M-128", "K4M-192"])
def test_second_order_one_million_cells_against_the_oracle(tile):
    """128-cell tiles run K3M (run-time shapes, gradient pass), 192-cell tiles K4M, the bench's second-order kernel (compile-time
    shape, gradients inside the tile kernel over a two-ring halo); with many tiles per CTA both pipeline stages and parities turn over."""
    case = _case(second_order=True)
    u_ref, _ = _reference_state("second_order", case)
    op = Operator.from_case(case)
    op.set_kernel(TILED, tile)
    op.step_euler(case.dt, STEPS)
    u = op.get_state()
    info = op.kernel_info()
    assert ("K4M" in info and "in-kernel gradients" in info) if tile == 192 else ("K3M" in info)
    assert_state_close(u, u_ref, what="1 M cells second order")
    op.close()


@pytest.mark.parametrize("hr", [False, True], ids=["first_order", "hr"])
def test_k4_one_million_cells_bitwise_repeatable_and_mass_conserving(hr):
    """No float atomics and a fixed accumulation order: the full grid, a 3-CTA grid and a repeat run give the same bits (which
    CTA processes a tile cannot matter); with closed walls and no rain the mass is conserved to round-off."""
    case = _case(hr=hr)
    case.sources[:] = 0.0
    runs = []
    for cap in (0, 0, 3):
        op = Operator.from_case(case)
        op.set_kernel(TILED_FIXED, 224)
        if cap:
            op.set_max_ctas(cap)
        m0 = op.mass()
        op.step_euler(case.dt, STEPS)
        runs.append(op.get_state())
        m1 = op.mass()
        op.close()
        assert abs(m1 - m0) <= 1e-12 * m0 * np.sqrt(STEPS)
    assert np.array_equal(runs[0], runs[1])
    assert np.array_equal(runs[0], runs[2])


@pytest.mark.parametrize("source_method", [cases.SEMI_IMPLICIT, cases.IMPLICIT_XQ2018])
@pytest.mark.parametrize("kernel", [(0, 0), (TILED_FIXED, 224)], ids=["K1", "K4"])
def test_subnormal_squared_speed_does_not_poison_the_state(kernel, source_method):
    """a wet cell whose squared speed u*u + v*v is subnormal (|u| ~ 1e-162..1e-154) made rsqrt.approx.ftz return
    Inf and the friction term NaN, which then spread over the mesh.  On axis-aligned quads the transverse momentum has no O(1)
    pressure term to be rounded against (sn == 0 exactly), so such values do occur ahead of a disturbance.  Here they are set
    directly: a uniform lake with |hu|, |hv| between 1e-163 and 1e-153 next to a moving patch; the reference's sqrt() returns a
    tiny finite value, the CUDA path 0 below 2^-1000 (swe_math.cuh) - both give a finite state that agrees to round-off."""
    from rdycore_b200.mesh import HostMesh

    mesh = HostMesh.quads(96, 64, 96.0, 64.0)
    bnd = cases.boundaries_from_predicates(mesh, [])
    n = mesh.num_owned_cells
    xc, yc = mesh.array("cell_xc"), mesh.array("cell_yc")
    rng = np.random.default_rng(42)
    u0 = np.zeros((n, 3))
    u0[:, 0] = 1.0
    u0[:, 1] = 10.0 ** rng.uniform(-163, -153, n) * rng.choice([-1.0, 1.0], n)
    u0[:, 2] = 10.0 ** rng.uniform(-163, -153, n) * rng.choice([-1.0, 1.0], n)
    patch = (xc < 8.0) & (yc < 8.0)
    u0[patch, 0] = 1.4
    u0[patch, 1] = 0.3
    case = cases.Case("subnormal_speeds", mesh, cases.make_config(source_method=source_method), bnd, u0, np.full(n, 0.03), None, 0.05)
    oracle = make_oracle(case)
    op = Operator.from_case(case)
    if kernel[0]:
        op.set_kernel(*kernel)
    for _ in range(4):
        op.step_euler(case.dt, 25)
        oracle.step_euler(case.dt, 25)
        u, u_ref = op.get_state(), oracle.array("u_global").reshape(-1, 3)
        assert not np.isnan(u_ref).any()
        assert not np.isnan(u).any(), "a subnormal squared speed produced NaN"
        assert_state_close(u, u_ref, what="subnormal speeds")
    op.close()
