"""The reference's own SWE decks on the reference's own fixture files: driver/tests/swe_roe/{ex2b_ic_file,
mixed_elements_ic_file,four_mounds_60x24,levee.hr,Houston1km.DirichletBC}.yaml with their Exodus meshes, PETSc-binary initial
conditions, Manning files and cell elevations (tests/golden/ref_fixtures/, tests/decks/ref/, copied by
tests/golden/make_ref_fixtures.py).  The files go through the C readers of include/rdyb200_io.h, the physics / numerics / time sections of
the decks (values unchanged) through RDyCreate; the list sections (regions, materials, conditions) are applied here through the setters.

CPU: restated oracle == the reference's own code, bitwise, on these meshes.  GPU: CUDA path == oracle at the stated tolerance
after the deck's own number of steps."""
import os

import numpy as np
import pytest

from rdycore_b200 import cases, io, rdy
from rdycore_b200.operator import GATHER, TILED_FIXED

from oracle import binding
from util import assert_state_close, celerity_momentum, make_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "ref_fixtures")
# deck -> files named by its grid / materials / flow_conditions sections, its time step in seconds and its number of steps
DECKS = {
    "ex2b_ic_file": dict(mesh="DamBreak_grid5x10", ic="DamBreak_grid5x10_wetdownstream.ic", manning="manning_grid5x10", dt=18.0 / 1000, steps=1000),
    "mixed_elements_ic_file": dict(mesh="DamBreak_grid5x10_mixed_elements", ic="DamBreak_grid5x10_mixed_elements_wetdownstream.ic",
                                   manning="manning_grid5x10_mixed_elements", dt=18.0 / 1000, steps=1000),
    "four_mounds_60x24": dict(mesh="four_mounds_60x24", ic="four_mounds_60x24.ic", manning="manning_four_mound_60x24", dt=0.02, steps=25),
    "levee.hr": dict(mesh="levee", ic="levee_ic", manning=0.033, elevation="levee_elevation", hr=True, dt=0.1, steps=600),
    "Houston1km.DirichletBC": dict(mesh="Houston1km_with_z", ic="Houston1km.ic", manning=0.015, dirichlet_zero=[1], dt=30.0, steps=140),
}


def _vec(stem):
    return io.read_petsc_vec(os.path.join(FIX, stem + ".int32.bin"))


def deck_case(name) -> cases.Case:
    d = DECKS[name]
    ex = io.read_exodus(os.path.join(FIX, d["mesh"] + ".exo"))
    mesh = io.mesh_from_exodus(ex)
    n = mesh.num_owned_cells
    if d.get("elevation"):  # grid.cell_elevation.file: zc of the hydrostatic reconstruction (src/swe/swe_petsc.c:1208-1224)
        zc = _vec(d["elevation"])
        assert zc.shape == (n,)
        mesh.set_zc(zc)
    bc = {sid: cases.DIRICHLET for sid in d.get("dirichlet_zero", [])}
    bnd = io.boundaries_from_side_sets(mesh, ex, bc_types=bc)
    u0 = _vec(d["ic"]).reshape(-1, 3)  # natural (file element) order, block size 3
    assert u0.shape == (n, 3)
    mann = np.full(n, d["manning"]) if isinstance(d["manning"], float) else _vec(d["manning"])
    assert mann.shape == (n,)
    cfg = cases.make_config(well_balancing=cases.WB_HR if d.get("hr") else cases.WB_NONE)
    case = cases.Case(name, mesh, cfg, bnd, u0, mann, None, d["dt"])
    for i, b in enumerate(bnd):
        if b.bc_type == cases.DIRICHLET:
            case.dirichlet[i] = np.zeros((len(b.edge_ids), 3))  # flow_conditions.outflow_bc: height, x_momentum, y_momentum = 0
    return case


@pytest.mark.parametrize("name", sorted(DECKS))
def test_port_equals_reference_on_the_reference_decks(name):
    if not binding.available("reference"):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    case = deck_case(name)
    ref, port = make_oracle(case, "reference"), make_oracle(case, "port")
    steps = min(DECKS[name]["steps"], 300)
    ref.step_euler(case.dt, steps)
    port.step_euler(case.dt, steps)
    u = ref.array("u_global")
    assert not np.isnan(u).any()
    assert np.array_equal(u, port.array("u_global"))
    assert np.abs(u.reshape(-1, 3) - case.u0).max() > 1e-6  # the state moved: not a static comparison
    assert ref.courant() < 1.0


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [(GATHER, 0), (TILED_FIXED, 224), (TILED_FIXED, 256), (TILED_FIXED, 224, 3)],
                         ids=["K1", "K4-224", "K4-256", "K4-224-cap3"])
@pytest.mark.parametrize("name", sorted(DECKS))
def test_cuda_path_runs_the_reference_decks(name, kernel, monkeypatch):
    """RDyCreate(<the deck's physics / numerics / time sections>) -> RDySetup -> RDyAdvance until RDyFinished, against the oracle."""
    if len(kernel) == 3:  # persistent grid capped at 3 CTAs: every CTA walks many tiles (the bench regime), via the process-wide default
        monkeypatch.setenv("RDYB_MAX_CTAS", str(kernel[2]))
        kernel = kernel[:2]
    case = deck_case(name)
    d = DECKS[name]
    oracle = make_oracle(case)
    oracle.step_euler(case.dt, d["steps"])
    rdy.RDyInit()
    r = rdy.RDy(os.path.join(HERE, "decks", "ref", name + ".yaml"))
    arr, nb = case.boundary_views()
    r.set_mesh(case.mesh, arr, nb)
    if kernel[0] != GATHER:
        r.set_kernel(*kernel)
    r.RDySetup()
    assert r.RDyGetTimeStep() == pytest.approx(case.dt, rel=1e-15)
    r.RDySetDomainManningsN(case.mannings)
    r.RDySetInitialConditions(case.u0)
    for i, vals in case.dirichlet.items():
        r.RDySetFlowDirichletBoundaryValues(i, len(vals), 3, vals)
    calls = 0
    while not r.RDyFinished():
        r.RDyAdvance()
        calls += 1
        assert calls <= d["steps"]
    assert r.RDyGetStep() == d["steps"]
    u = r.state()
    r.RDyDestroy()
    # ex2b_ic_file and levee.hr end (almost) at rest — |momentum| <= 0.008 under 1.5 to 7 m of water — so a momentum error relative to
    # the momentum itself only measures the round-off of the pressure fluxes: the same C source with and without FMA contraction
    # differs by 2.9e-10 / 7e-11 on those two decks (scratch/fma/sens4.py).  Momentum is therefore compared against
    # max(|q|max, h sqrt(g h)); depth keeps the plain relative bar.
    u_ref = oracle.array("u_global")
    assert_state_close(u, u_ref, what=name, qscale_floor=celerity_momentum(u_ref))
