"""GPU: the B200 flux / source sub-operators, driven exactly the way ApplyPetscOperator drives the reference's
(src/operator.c:656-672), against the CPU oracle (the reference's own swe_petsc.c where available)."""
import numpy as np
import pytest

from rdycore_b200 import cases
from rdycore_b200 import petsc_operator as po
from rdycore_b200._lib import RdybCourantDiagnostics, RdybError
from rdycore_b200.mesh import HostMesh

from util import make_oracle

pytestmark = pytest.mark.gpu


class RefStyleOperator:
    """The PETSc half of the reference's Operator (src/operator.c:348-417, 656-672) on top of the B200 sub-operators."""

    def __init__(self, case, fused=False, with_fields=True):
        self.case = case
        self.arr, self.nb = case.boundary_views()
        self.view = case.mesh.view()
        n, nl = self.view.num_owned_cells, self.view.num_cells
        self.n = n
        ne = [len(b.edge_ids) for b in case.boundaries]
        # src/operator.c:368-379: sequential Vecs with block size 3, one block per boundary edge
        self.boundary_values = [po.Vec(3 * e, 3) for e in ne]
        self.boundary_fluxes = [po.Vec(3 * e, 3) for e in ne]
        self.boundary_fluxes_accum = [po.Vec(3 * e, 3) for e in ne]
        self.external_sources, self.material_properties = po.Vec(3 * n, 3), po.Vec(n, 1)
        self.diagnostics = RdybCourantDiagnostics(0.0, -1, -1)
        self.flux = po.CreateB200FluxOperator(case.config, self.view, self.arr, self.nb, self.boundary_values, self.boundary_fluxes,
                                              self.boundary_fluxes_accum, self.diagnostics)
        self.source = po.CreateB200SourceOperator(case.config, self.view, self.external_sources, self.material_properties, self.flux)
        po.set_fused(self.flux, fused)
        self.flux_divergence, self.primitive_variables = po.Vec(3 * n, 3), po.Vec(3 * n, 3)
        if with_fields:  # AddOperatorFluxDivergence / AddOperatorPrimitiveVariables, src/operator.c:224-299
            po.PetscOperatorSetField(self.source, "riemannf", self.flux_divergence)
            po.PetscOperatorSetField(self.source, "primitive_variables", self.primitive_variables)
        self.u_local, self.F = po.Vec(3 * nl, 3, pinned=True), po.Vec(3 * n, 3, pinned=True)
        self.material_properties.set(case.mannings)
        if case.sources is not None:
            self.external_sources.set(case.sources)
        for b, vals in case.dirichlet.items():
            self.boundary_values[b].set(vals)

    def apply(self, dt, u_local):
        """OperatorRHSFunction + ApplyPetscOperator: F <- 0; flux; VecCopy(F, flux_divergence); source."""
        self.u_local.set(u_local)
        self.F.zero()
        self.diagnostics.max_courant_num = 0.0  # ResetOperatorDiagnostics, src/rdysetup.c:1136
        po.PetscOperatorApply(self.flux, dt, self.u_local, self.F)
        self.F.copy_to(self.flux_divergence)
        po.PetscOperatorApply(self.source, dt, self.u_local, self.F)
        return self.F.array().reshape(-1, 3).copy()

    def destroy(self):
        po.PetscOperatorDestroy(self.source)
        po.PetscOperatorDestroy(self.flux)


def _close(a, b, tol=1e-13):
    a, b = np.asarray(a).reshape(-1, 3), np.asarray(b).reshape(-1, 3)
    scale = np.maximum(np.abs(b).max(axis=0), 1e-300)
    return (np.abs(a - b) / scale).max() < tol


def _moving_floodplain(hr=False, source_method=cases.SEMI_IMPLICIT):
    case = cases.floodplain(nx=40, ny=48, eta0=9.0, hr=hr, source_method=source_method)
    rng = np.random.default_rng(11)
    case.u0[:, 1] = rng.normal(0, 0.3, case.u0.shape[0]) * case.u0[:, 0]
    case.u0[:, 2] = rng.normal(0, 0.3, case.u0.shape[0]) * case.u0[:, 0]
    case.sources[:, 1] = 1e-4 * rng.normal(size=case.u0.shape[0])
    case.sources[:, 2] = 1e-4 * rng.normal(size=case.u0.shape[0])
    return case


@pytest.mark.parametrize("hr", [False, True])
@pytest.mark.parametrize("source_method", [cases.SEMI_IMPLICIT, cases.IMPLICIT_XQ2018])
def test_split_flux_then_source_matches_the_reference_at_every_point_of_the_abi(hr, source_method):
    case = _moving_floodplain(hr, source_method)
    oracle = make_oracle(case)
    F_ref = oracle.rhs(case.dt).reshape(-1, 3).copy()
    op = RefStyleOperator(case)
    for _ in range(2):  # the second apply takes the steady path (primitive variables produced by the one kernel run)
        F = op.apply(case.dt, oracle.array("u_local"))
        assert _close(F, F_ref)
        # after flux.apply F held the flux divergence: that is what the reference's VecCopy put into "riemannf"
        assert _close(op.flux_divergence.array(), oracle.array("flux_divergence"))
        assert np.abs(op.primitive_variables.array() - oracle.array("primitive_variables")).max() < 1e-13
        assert op.diagnostics.max_courant_num == pytest.approx(oracle.courant(), rel=1e-12)
    op.destroy()


def test_fused_mode_writes_the_complete_rhs_in_flux_apply():
    case = _moving_floodplain()
    oracle = make_oracle(case)
    F_ref = oracle.rhs(case.dt).reshape(-1, 3).copy()
    op = RefStyleOperator(case, fused=True, with_fields=False)
    op.u_local.set(oracle.array("u_local"))
    op.F.zero()
    po.PetscOperatorApply(op.flux, case.dt, op.u_local, op.F)
    assert _close(op.F.array(), F_ref)
    po.PetscOperatorApply(op.source, case.dt, op.u_local, op.F)  # no-op
    assert _close(op.F.array(), F_ref)
    op.destroy()


def test_changed_host_vecs_are_picked_up_by_the_next_apply():
    """Restore*ExternalSource / *MaterialProperties / SetOperatorBoundaryValues write host Vecs between applies
    (src/operator.c:1045-1061, 1394-1464): the next apply must see them."""
    mesh = HostMesh.quads(24, 12, 12.0, 6.0)
    bnd = cases.boundaries_from_predicates(mesh, [
        (1, "west", cases.DIRICHLET, lambda x, y: x < 1e-9),
        (2, "east", cases.CRITICAL_OUTFLOW, lambda x, y: x > 12.0 - 1e-9),
    ])
    n = mesh.num_owned_cells
    u0 = np.zeros((n, 3))
    u0[:, 0] = 1.0 + 0.1 * np.sin(np.arange(n))
    u0[:, 1] = 0.2
    ne = len(bnd[0].edge_ids)
    dvals = np.tile(np.array([1.5, 0.75, 0.0]), (ne, 1))
    case = cases.Case("channel", mesh, cases.make_config(), bnd, u0, np.full(n, 0.03), np.zeros((n, 3)), 2e-3, dirichlet={0: dvals})
    oracle = make_oracle(case)
    op = RefStyleOperator(case)
    F_ref = oracle.rhs(case.dt).reshape(-1, 3).copy()
    assert _close(op.apply(case.dt, oracle.array("u_local")), F_ref)
    for b in range(len(bnd)):  # boundary_fluxes of this apply and the dt-weighted accumulation (src/swe/swe_petsc.c:574-623)
        ref = oracle.array("boundary_fluxes", b)
        ok = ~np.isnan(ref)
        assert np.abs(op.boundary_fluxes[b].array()[ok] - ref[ok]).max() <= 1e-12 * max(np.abs(ref[ok]).max(), 1e-30)
        acc = oracle.array("boundary_fluxes_accum", b)
        ok = ~np.isnan(acc)
        assert np.abs(op.boundary_fluxes_accum[b].array()[ok] - acc[ok]).max() <= 1e-12 * max(np.abs(acc[ok]).max(), 1e-30)
    # change everything a caller may change
    rng = np.random.default_rng(3)
    new_src = np.zeros((n, 3))
    new_src[:, 0] = 1e-4 * rng.random(n)
    new_n = 0.02 + 0.03 * rng.random(n)
    new_d = dvals * 1.2
    oracle.array("external_sources")[:] = new_src.ravel()
    oracle.array("material_properties")[:] = new_n
    oracle.array("boundary_values", 0)[:] = new_d.ravel()
    op.external_sources.set(new_src)
    op.material_properties.set(new_n)
    op.boundary_values[0].set(new_d)
    F_ref2 = oracle.rhs(case.dt).reshape(-1, 3).copy()
    assert not _close(F_ref2, F_ref)
    assert _close(op.apply(case.dt, oracle.array("u_local")), F_ref2)
    op.destroy()


def test_euler_loop_through_the_sub_operators_matches_the_oracle_on_config_a():
    """TSEULER around the split operators: what RDyAdvance does with the PETSc backend (100x100 dam break would take
    minutes through host Vecs; 40x40 for 200 steps exercises the same code)."""
    case = cases.dam_break_quads(nx=40, ny=40)
    oracle = make_oracle(case)
    op = RefStyleOperator(case)
    u = case.u0.copy()
    for _ in range(200):
        u += case.dt * op.apply(case.dt, u)  # serial: local == owned order
    oracle.step_euler(case.dt, 200)
    u_ref = oracle.array("u_global").reshape(-1, 3)
    assert (np.abs(u - u_ref) / np.abs(u_ref).max(axis=0).clip(1e-300)).max() < 1e-12
    op.destroy()


def test_error_paths():
    case = cases.dam_break_quads(nx=8, ny=8)
    op = RefStyleOperator(case)
    n = op.n
    # block size of the local vector must be 3 (src/swe/swe_petsc.c:231-233)
    bad = po.Vec(3 * n, 1)
    with pytest.raises(RdybError) as ei:
        po.PetscOperatorApply(op.flux, 1e-3, bad, op.F)
    assert ei.value.code == 83 and "Number of dof in local vector must be 3" in str(ei.value)
    # the source operator hands over the result of the flux operator's run: applying it first is an error
    with pytest.raises(RdybError):
        po.PetscOperatorApply(op.source, 1e-3, op.u_local, op.F)
    # wrong sizes
    with pytest.raises(RdybError):
        po.PetscOperatorApply(op.flux, 1e-3, po.Vec(3 * (n + 1), 3), op.F)
    # destroying in either order is fine (the GPU operator is shared and reference-counted)
    po.PetscOperatorDestroy(op.flux)
    po.PetscOperatorDestroy(op.source)
