"""CPU: the oracles (test infrastructure) are pinned before anything is compared against them.

  1. golden vectors: tests/golden/swe_golden.npz was recorded by tests/golden/make_golden.py from oracle/_ref, i.e. the
     REFERENCE's own src/swe/swe_petsc.c + swe_roe_flux_petsc.h + petsc_operator.c + operator_{fluxes,sources}_petsc.c compiled
     where they lie; the port (oracle/swe_oracle.c, the from-scratch restatement that always travels) must reproduce them
     bit for bit, and so must oracle/_ref itself wherever it is present;
  2. port == reference bit for bit on further seeded cases (when oracle/_ref is available);
  3. known answers that hold for ANY correct implementation of the path: lake at rest on a flat bed gives F = 0, a uniform
     stream through a straight reflecting channel gives zero mass flux, closed basins conserve mass to round-off, the Roe
     flux of two equal states is the physical flux (consistency), reflecting walls mirror exactly.
The reference repository has no flux / source known-answer vectors of its own; its quantitative pins for this
path are the MMS convergence rates, covered in tests/test_mms.py.
"""
import os

import numpy as np
import pytest

from golden_cases import GOLDEN, run_oracle
from oracle import binding
from rdycore_b200 import cases
from rdycore_b200.mesh import HostMesh
from util import make_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN_PATH = os.path.join(HERE, "golden", "swe_golden.npz")
KINDS = ["port"] + (["reference"] if binding.available("reference") or os.path.isdir("/root/reference/src/swe") else [])


@pytest.fixture(scope="module")
def golden():
    assert os.path.exists(GOLDEN_PATH), "tests/golden/swe_golden.npz is missing (python tests/golden/make_golden.py)"
    return np.load(GOLDEN_PATH)


def _same_bits(a, b):
    a, b = np.asarray(a), np.asarray(b)
    return a.shape == b.shape and np.array_equal(a.view(np.uint64), b.view(np.uint64))


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("name", sorted(GOLDEN))
def test_oracle_reproduces_golden_vectors_bit_for_bit(golden, name, kind):
    builder, mode, nsteps = GOLDEN[name]
    case = builder()
    oracle = make_oracle(case, kind=kind)
    rec = run_oracle(oracle, case, mode, nsteps)
    keys = [k.split("/", 1)[1] for k in golden.files if k.startswith(name + "/") and not k.endswith("/digest")]
    assert keys, f"no golden record for {name}"
    for k in keys:
        want, got = golden[f"{name}/{k}"], rec[k]
        assert _same_bits(got, want), f"{kind} oracle differs from the golden vector {name}/{k}: max |d| = {np.nanmax(np.abs(got - want)):.3e}"


@pytest.mark.skipif("reference" not in KINDS, reason="oracle/_ref is not available here")
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_port_equals_reference_on_random_states(seed):
    """One RHS on a random (partly dry, partly supercritical) state, both friction methods and HR: every output Vec identical."""
    rng = np.random.default_rng(seed)
    for hr in (False, True):
        for sm in (cases.SEMI_IMPLICIT, cases.IMPLICIT_XQ2018):
            case = cases.floodplain(nx=12, ny=10, eta0=7.0, hr=hr, source_method=sm)
            n = case.u0.shape[0]
            case.u0[:, 0] *= rng.random(n) > 0.2  # carve dry holes
            case.u0[:, 1] = rng.normal(0, 3.0, n) * case.u0[:, 0]
            case.u0[:, 2] = rng.normal(0, 3.0, n) * case.u0[:, 0]
            case.sources[:, 1:] = 1e-3 * rng.normal(size=(n, 2))
            outs = []
            for kind in ("port", "reference"):
                o = make_oracle(case, kind=kind)
                rhs = o.rhs(case.dt).copy()
                outs.append((rhs, o.array("flux_divergence").copy(), o.array("primitive_variables").copy(), o.courant()))
            for a, b in zip(outs[0][:3], outs[1][:3]):
                assert _same_bits(a, b)
            assert outs[0][3] == outs[1][3]


@pytest.mark.parametrize("kind", KINDS)
def test_lake_at_rest_on_a_flat_bed_has_zero_rhs(kind):
    case = cases.dam_break_quads(12, 8)
    case.u0[:, 0] = 3.0
    o = make_oracle(case, kind=kind)
    rhs = o.rhs(case.dt).reshape(-1, 3)
    # the critical-outflow wall (y = 0) drains even a resting lake; every cell away from it stays exactly at rest
    yc = case.mesh.array("cell_yc")
    away = yc > yc.min() + 1e-9
    assert np.abs(rhs[away]).max() <= 1e-12
    assert rhs[~away, 0].max() < 0.0


@pytest.mark.parametrize("kind", KINDS)
def test_closed_basin_conserves_mass_to_round_off(kind):
    case = cases.floodplain(nx=16, ny=12, eta0=8.0)
    case.sources[:] = 0.0
    area = case.mesh.view_arrays()["cell_areas"][case.mesh.array("owned_to_local")]
    o = make_oracle(case, kind=kind)
    m0 = float((area * o.array("u_global").reshape(-1, 3)[:, 0]).sum())
    o.step_euler(case.dt, 100)
    m1 = float((area * o.array("u_global").reshape(-1, 3)[:, 0]).sum())
    assert abs(m1 - m0) <= 1e-12 * m0 * 10


@pytest.mark.parametrize("kind", KINDS)
def test_roe_flux_is_consistent_and_reflecting_wall_mirrors(kind):
    """Uniform stream along a straight channel with reflecting side walls: every interior cell sees equal left/right states,
    so the Roe flux reduces to the physical flux and its divergence vanishes; the walls add only pressure (no mass, no
    x-momentum).  h, u chosen so that the flow is subcritical."""
    mesh = HostMesh.quads(10, 4, 10.0, 4.0)
    bnd = cases.boundaries_from_predicates(mesh, [
        (1, "west", cases.DIRICHLET, lambda x, y: x < 1e-9),
        (2, "east", cases.DIRICHLET, lambda x, y: x > 10.0 - 1e-9),
    ])
    n = mesh.num_owned_cells
    h, u = 2.0, 1.5
    u0 = np.tile(np.array([h, h * u, 0.0]), (n, 1))
    dv = {0: np.tile(np.array([h, h * u, 0.0]), (len(bnd[0].edge_ids), 1)), 1: np.tile(np.array([h, h * u, 0.0]), (len(bnd[1].edge_ids), 1))}
    case = cases.Case("stream", mesh, cases.make_config(), bnd, u0, np.zeros(n), None, 1e-3, dirichlet=dv)
    o = make_oracle(case, kind=kind)
    rhs = o.rhs(case.dt).reshape(-1, 3)
    assert np.abs(rhs).max() <= 1e-12 * (h * u * u + 0.5 * cases.GRAVITY * h * h)
    # Courant number of the uniform stream: (|u| + sqrt(g h)) * L / A * dt with L / A = 1 for unit squares
    assert o.courant() == pytest.approx((u + np.sqrt(cases.GRAVITY * h)) * case.dt, rel=1e-12)


@pytest.mark.parametrize("reconstruction", [cases.MUSCL_MINMOD, cases.MUSCL_NONE, cases.MUSCL_VANLEER])
def test_port_equals_reference_on_the_second_order_path(reconstruction):
    """ApplyInteriorFlux2R compiled from the reference tree (over the restated helpers of oracle/muscl_helpers.h) against the
    port's restatement of the loop: bit for bit, on quads and on jittered triangles."""
    if not (binding.available("reference") and binding.available("port")):
        pytest.skip("needs both oracles")
    def wet_floodplain():
        c = cases.floodplain(nx=24, ny=20, eta0=40.0)
        rng = np.random.default_rng(1)
        c.u0[:, 1] = 0.05 * rng.normal(size=len(c.u0)) * c.u0[:, 0]
        return c
    makers = [lambda: cases.dam_break_quads(nx=30, ny=30, source_method=cases.SEMI_IMPLICIT)]
    if reconstruction != cases.MUSCL_NONE:  # unlimited extrapolation over the rough synthetic bathymetry blows up (in the reference too)
        makers.append(wet_floodplain)
    for make in makers:
        res = []
        for kind in ("reference", "port"):
            case = make()
            case.config.reconstruction = reconstruction
            o = make_oracle(case, kind)
            o.step_euler(case.dt, 60)
            res.append((o.array("u_global").copy(), o.courant()))
        assert not np.isnan(res[0][0]).any()
        assert _same_bits(res[0][0], res[1][0]) and res[0][1] == res[1][1]
