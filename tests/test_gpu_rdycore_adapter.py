"""The drop-in boundary with the REFERENCE'S OWN factory signatures.

rdycore_b200/csrc/rdycore_adapter.c is compiled against the reference's headers (include/private/rdycoreimpl.h,
rdyoperatorimpl.h; behind oracle/ref_shim here, behind PETSc in an RDycore build) and exports CreateB200FluxOperator /
CreateB200FluxHROperator / CreateB200SourceOperator / CreateB200SourceHROperator with exactly the argument lists of
CreatePetscFlux[HR]Operator / CreatePetscSource[HR]Operator (include/private/rdyoperatorimpl.h:234-239), taking RDyConfig*,
RDyMesh*, RDyBoundary*, RDyCondition*, the Operator's Vecs and OperatorDiagnostics* as they are.

oracle/_ref/libswe_adapter_ref.so (oracle/Makefile target `adapter`) is the reference driver of the oracle with ONLY those four
calls switched: the Operator struct, its Vecs, the reference's src/petsc_operator.c (PetscOperatorCreate / Composite / SetField /
Apply) and the four statements of ApplyPetscOperator (src/operator.c:656-672) are the same code in both libraries.  So the same
inputs go once through the reference's sub-operators and once through the B200 ones, and every Vec the reference fills is compared:
F, "riemannf" (flux_divergence), "primitive_variables", boundary_fluxes / boundary_fluxes_accum, the Courant diagnostic.
"""
import numpy as np
import pytest

from oracle import binding
from rdycore_b200 import cases
from rdycore_b200.mesh import HostMesh

from util import assert_state_close, make_oracle

needs_adapter = pytest.mark.skipif(not (binding.available("adapter") and binding.available("reference")),
                                   reason="oracle/_ref/libswe_adapter_ref.so not built (needs /root/reference at build time)")


def _moving(case, seed=7):
    rng = np.random.default_rng(seed)
    case.u0[:, 1] = rng.normal(0, 0.3, case.u0.shape[0]) * case.u0[:, 0]
    case.u0[:, 2] = rng.normal(0, 0.3, case.u0.shape[0]) * case.u0[:, 0]
    return case


def _dirichlet_channel():
    mesh = HostMesh.quads(40, 20, 20.0, 10.0)
    bnd = cases.boundaries_from_predicates(mesh, [
        (1, "west", cases.DIRICHLET, lambda x, y: x < 1e-9),
        (2, "east", cases.CRITICAL_OUTFLOW, lambda x, y: x > 20.0 - 1e-9),
    ])
    n = mesh.num_owned_cells
    u0 = np.zeros((n, 3))
    u0[:, 0] = 1.0
    u0[:, 1] = 0.4
    ne = len(bnd[0].edge_ids)
    dvals = np.tile(np.array([1.5, 0.75, 0.0]), (ne, 1)) * (1 + 0.1 * np.arange(ne)[:, None] / ne)
    return cases.Case("dirichlet_channel", mesh, cases.make_config(), bnd, u0, np.full(n, 0.03), None, 2e-3, dirichlet={0: dvals})


CASES = {
    "floodplain_semi_implicit": lambda: _moving(cases.floodplain(nx=40, ny=48, eta0=9.0)),
    "floodplain_xq2018": lambda: _moving(cases.floodplain(nx=40, ny=48, eta0=9.0, source_method=cases.IMPLICIT_XQ2018)),
    "floodplain_hr": lambda: _moving(cases.floodplain(nx=48, ny=40, eta0=6.0, hr=True, dam_column=(150.0, 12.0))),
    "dam_break_quads_outflow": lambda: _moving(cases.dam_break_quads(60, 40), seed=3),
    "dirichlet_channel": _dirichlet_channel,
}


@pytest.mark.gpu
@needs_adapter
@pytest.mark.parametrize("name", sorted(CASES))
def test_apply_petsc_operator_with_reference_and_b200_suboperators(name):
    case = CASES[name]()
    if case.sources is not None:  # 3-component external sources: the source operator's S_h, S_hu, S_hv
        rng = np.random.default_rng(11)
        case.sources[:, 1] = 1e-4 * rng.normal(size=case.u0.shape[0])
        case.sources[:, 2] = 1e-4 * rng.normal(size=case.u0.shape[0])
    ref = make_oracle(case, "reference")
    b200 = make_oracle(case, "adapter")  # the same driver, sub-operators made by CreateB200*Operator
    assert b200.lib.swe_oracle_kind().decode() == "adapter"
    for _ in range(2):  # twice: accumulated boundary fluxes, re-used device operator
        F_ref = ref.rhs(case.dt).reshape(-1, 3).copy()
        F = b200.rhs(case.dt).reshape(-1, 3).copy()
        scale = np.maximum(np.abs(F_ref).max(axis=0), 1e-300)  # (a channel flow has an identically zero y-momentum RHS)
        assert (np.abs(F - F_ref) / scale).max() <= 1e-13
        fd_ref, fd = ref.array("flux_divergence").reshape(-1, 3), b200.array("flux_divergence").reshape(-1, 3)
        assert np.abs(fd_ref).max() > 0
        assert (np.abs(fd - fd_ref) / np.maximum(np.abs(fd_ref).max(axis=0), 1e-300)).max() <= 1e-13
        assert np.abs(b200.array("primitive_variables") - ref.array("primitive_variables")).max() <= 1e-13 * max(1.0, np.abs(ref.array("primitive_variables")).max())
        assert np.array_equal(b200.array("src_inst"), ref.array("src_inst"))
        for b, bd in enumerate(case.boundaries):
            if not len(bd.edge_ids):
                continue
            for vec in ("boundary_fluxes", "boundary_fluxes_accum"):
                a_ref, a = ref.array(vec, b), b200.array(vec, b)
                ok = ~np.isnan(a_ref)  # the reference leaves NaN on dry-dry boundary edges; the CUDA path stores 0
                if ok.any():
                    assert np.abs(a[ok] - a_ref[ok]).max() <= 1e-13 * max(np.abs(a_ref[ok]).max(), 1e-30), vec
        assert b200.courant() == pytest.approx(ref.courant(), rel=1e-12)


@pytest.mark.gpu
@needs_adapter
@pytest.mark.parametrize("name", ["floodplain_semi_implicit", "floodplain_hr", "dam_break_quads_outflow"])
def test_time_stepping_through_the_reference_signature_factories(name):
    """TSEULER driven from the host (the Level-1 integration: host Vecs every stage) over the B200 sub-operators."""
    case = CASES[name]()
    ref = make_oracle(case, "reference")
    b200 = make_oracle(case, "adapter")
    ref.step_euler(case.dt, 60)
    b200.step_euler(case.dt, 60)
    assert_state_close(b200.array("u_global"), ref.array("u_global"), what=f"adapter {name}")


@needs_adapter
def test_adapter_library_exports_the_reference_signature_factories_and_fails_loudly_without_a_gpu():
    import ctypes as C

    lib = binding.load("adapter")
    for sym in ("CreateB200FluxOperator", "CreateB200FluxHROperator", "CreateB200SourceOperator", "CreateB200SourceHROperator",
                "RDyB200GetDeviceOperator", "RDyB200SetFused"):
        assert hasattr(lib, sym), sym
    # none of the reference's flux / source code is linked into it: the sub-operators can only be the B200 ones
    for sym in ("CreatePetscFluxOperator", "CreatePetscSWEInteriorFluxOperator", "CreatePetscSourceOperator"):
        assert not hasattr(lib, sym), sym
    try:
        import torch

        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        case = cases.floodplain(nx=8, ny=8, eta0=9.0)
        arr, nb = case.boundary_views()
        view = case.mesh.view()
        h = C.c_void_p()
        ierr = lib.swe_oracle_create(C.byref(case.config), C.byref(view), C.c_int32(nb), arr, C.byref(h))
        assert ierr == 97  # RDYB_ERR_CUDA: no CPU fallback behind the factories
