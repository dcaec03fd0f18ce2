"""RDycore's public lifecycle (include/rdycore_b200.h, mirroring include/rdycore.h) on the B200 path.

CPU part: deck reading and the reference's configuration checks (src/yaml_input.c:905-1000), argument checks.
GPU part: RDyCreate -> RDySetup -> while (!RDyFinished) RDyAdvance, compared with the oracle driven by a restatement
of RDyAdvance's time loop (src/rdyadvance.c:261-383).
"""
import math
import os

import numpy as np
import pytest

from rdycore_b200 import cases, rdy
from rdycore_b200._lib import RdybError

from util import assert_state_close, make_oracle

DECKS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "decks")


@pytest.fixture(autouse=True)
def _init():
    rdy.RDyInit()
    yield


def _deck(tmp_path, text):
    p = tmp_path / "deck.yaml"
    p.write_text(text)
    return str(p)


BASE = """
physics:
  flow:
    mode: swe
numerics:
  spatial: fv
  temporal: euler
  riemann: roe
"""


def test_create_reads_the_scalar_sections_and_skips_the_lists():
    r = rdy.RDy(os.path.join(DECKS, "dam_break.yaml"))
    assert r.RDyGetTimeUnit() == rdy.TIME_SECONDS
    assert r.RDyGetCouplingInterval() == pytest.approx(0.05)
    assert r.RDyGetStep() == 0 and r.RDyGetTime() == 0.0
    assert not r.RDyFinished()
    # units convert like RDyConvertTime (src/rdyadvance.c:420-443)
    assert r.RDyGetCouplingInterval(rdy.TIME_MINUTES) == pytest.approx(0.05 / 60)
    r.RDySetCouplingInterval(rdy.TIME_MINUTES, 0.002)
    assert r.RDyGetCouplingInterval() == pytest.approx(0.12)
    r.RDyDestroy()


def test_create_requires_init_and_an_existing_file(tmp_path):
    rdy.RDyFinalize()
    with pytest.raises(RdybError):
        rdy.RDy(os.path.join(DECKS, "dam_break.yaml"))
    rdy.RDyInit()
    with pytest.raises(RdybError) as ei:
        rdy.RDy(str(tmp_path / "missing.yaml"))
    assert "Could not open" in str(ei.value)


@pytest.mark.parametrize("time_section,message", [
    ("time:\n  unit: seconds\n  stop: 1.0\n", "Exactly 2 of time.final_time, time.max_step, time.time_step"),
    ("time:\n  unit: seconds\n  stop: 1.0\n  stop_n: 10\n  time_step: 0.1\n", "Exactly 2 of"),
    ("time:\n  stop: 1.0\n  stop_n: 10\n", "time.unit must be specified"),
    ("time:\n  unit: seconds\n  stop: 1.0\n  stop_n: 10\n  coupling_interval: 2.0\n", "must not exceed time.final_time"),
    ("time:\n  unit: seconds\n  stop: 1.0\n  stop_n: 10\n  coupling_interval: -1\n", "must be positive"),
    ("time:\n  unit: seconds\n  stop: 1.0\n  stop_n: 10\n  adaptive:\n    enable: true\n    target_courant_number: 0.5\n"
     "    max_increase_factor: 2\n    initial_time_step: 0.1\n", "max_step cannot be specified with adaptive"),
    ("time:\n  unit: seconds\n  stop: 1.0\n  adaptive:\n    enable: true\n    target_courant_number: 1.5\n"
     "    max_increase_factor: 2\n    initial_time_step: 0.1\n", "must be less than 1.0"),
    ("time:\n  unit: fortnights\n  stop: 1.0\n  stop_n: 10\n", "Invalid value for time.unit"),
])
def test_time_section_checks_match_the_reference(tmp_path, time_section, message):
    with pytest.raises(RdybError) as ei:
        rdy.RDy(_deck(tmp_path, BASE + time_section))
    assert ei.value.code == 83 and message in str(ei.value)


def test_physics_and_numerics_checks(tmp_path):
    t = "time:\n  unit: seconds\n  stop: 1.0\n  stop_n: 10\n"
    with pytest.raises(RdybError) as ei:  # src/yaml_input.c:910-912
        rdy.RDy(_deck(tmp_path, "physics:\n  flow:\n    mode: swe\n    source:\n      method: semi_implicit\n      xq2018_threshold: 1e-8\n" + t))
    assert "Erroneously set XQ2018 threshold" in str(ei.value)
    with pytest.raises(RdybError) as ei:
        rdy.RDy(_deck(tmp_path, "physics:\n  flow:\n    mode: swe\nnumerics:\n  temporal: beuler\n" + t))
    assert ei.value.code == 56
    with pytest.raises(RdybError):
        rdy.RDy(_deck(tmp_path, "physics:\n  flow:\n    mode: swe\nnumerics:\n  riemann: hllc\n" + t))
    with pytest.raises(RdybError):
        rdy.RDy(_deck(tmp_path, "physics:\n  flow:\n    mode: diffusion\n" + t))
    # third time parameter is derived from the two given (src/yaml_input.c:978-986)
    r = rdy.RDy(_deck(tmp_path, BASE + "time:\n  unit: minutes\n  stop: 2.0\n  stop_n: 100\n"))
    assert r.RDyGetCouplingInterval(rdy.TIME_MINUTES) == pytest.approx(2.0)  # default: the whole run
    r.RDyDestroy()


def test_setup_needs_a_mesh_and_a_gpu(tmp_path):
    import torch

    r = rdy.RDy(os.path.join(DECKS, "dam_break.yaml"))
    with pytest.raises(RdybError) as ei:
        r.RDySetup()
    assert "RDyB200SetMesh" in str(ei.value)
    with pytest.raises(RdybError):
        r.RDyAdvance()  # not set up
    case = cases.dam_break_quads(nx=8, ny=8)
    arr, nb = case.boundary_views()
    r.set_mesh(case.mesh, arr, nb)
    assert r.RDyGetNumOwnedCells() == 64 and r.RDyGetNumBoundaryConditions() == nb
    assert r.RDyGetBoundaryConditionFlowType(1) == cases.CRITICAL_OUTFLOW and r.RDyGetBoundaryID(0) == 2
    with pytest.raises(RdybError) as ei:  # src/rdydata.c:24-32
        r.RDyGetNumBoundaryEdges(nb)
    assert "exceeds the max number of boundary conditions" in str(ei.value)
    if not torch.cuda.is_available():
        with pytest.raises(RdybError) as ei:
            r.RDySetup()
        assert ei.value.code == 97 and "no CPU fallback" in str(ei.value)
    r.RDyDestroy()


# ---------------------------------------------------------------------------------------------------------------
# GPU
# ---------------------------------------------------------------------------------------------------------------
def _make(deck, case, regions=()):
    r = rdy.RDy(os.path.join(DECKS, deck))
    arr, nb = case.boundary_views()
    r.set_mesh(case.mesh, arr, nb)
    for ids in regions:
        r.add_region(ids)
    r.RDySetup()
    r.RDySetDomainManningsN(case.mannings)
    if case.sources is not None:
        r.RDySetDomainWaterSource(case.sources[:, 0])
    r.RDySetInitialConditions(case.u0)
    return r


@pytest.mark.gpu
def test_dam_break_through_the_public_lifecycle_matches_the_oracle():
    case = cases.dam_break_quads(nx=50, ny=50, dt=1e-3)
    r = _make("dam_break.yaml", case)
    oracle = make_oracle(case)
    n_adv = 0
    while not r.RDyFinished():
        r.RDyAdvance()
        n_adv += 1
    assert n_adv == 4 and r.RDyGetStep() == 200 and r.RDyGetTime() == pytest.approx(0.2)
    assert r.RDyGetTimeStep() == pytest.approx(1e-3)
    oracle.step_euler(1e-3, 200)
    assert_state_close(r.state(), oracle.array("u_global"), what="RDyAdvance dam break")
    # getters: sizes are checked like CheckNumOwnedCells (src/rdydata.c:54-58)
    with pytest.raises(RdybError):
        r.RDyGetOwnedCellHeights(size=7)
    xc = r.RDyGetOwnedCellXCentroids()
    assert np.allclose(xc, case.mesh.array("cell_xc")[: len(xc)])
    assert np.allclose(r.RDyGetOwnedCellAreas(), 10.0 / 50 * 5.0 / 50)
    r.RDyDestroy()


@pytest.mark.gpu
def test_rk4_with_a_shortened_last_step():
    """time_step 0.002 x 55 steps = 0.11 s, one coupling interval: 55 whole RK4 steps."""
    case = cases.dam_break_quads(nx=30, ny=30, source_method=cases.SEMI_IMPLICIT)
    case.config.temporal = cases.RK4
    r = _make("rk4.yaml", case)
    r.RDySetCouplingInterval(rdy.TIME_SECONDS, 0.025)  # 12.5 steps per interval: every interval ends with a half step
    oracle = make_oracle(case)
    t = 0.0
    while not r.RDyFinished():
        r.RDyAdvance()
        # oracle: the same TS loop
        t_end = t + 0.025  # RDyAdvance always covers a whole coupling interval (src/rdyadvance.c:296-297)
        while t < t_end - 1e-12:
            dt = min(0.002, t_end - t)
            oracle.step_rk4(dt, 1)
            t += dt
    assert r.RDyGetTime() == pytest.approx(t)
    assert_state_close(r.state(), oracle.array("u_global"), what="RDyAdvance rk4")
    r.RDyDestroy()


@pytest.mark.gpu
def test_adaptive_time_step_follows_the_reference_rule():
    """src/rdyadvance.c:303-343: dt is rescaled at the start of every RDyAdvance from the Courant number of the last
    right-hand-side evaluation of the previous one.  (That evaluation may be the shortened last step of the interval, whose
    small dt gives a small Courant number and lets dt grow beyond the target: reference behaviour, mirrored as it is.)"""
    # a fully wet case: with wet/dry fronts a 1e-15 relative change of dt moves the state by 1e-8 (measured with the oracle),
    # which would test chaos, not the time loop
    case = cases.dam_break_quads(nx=40, ny=40, source_method=cases.SEMI_IMPLICIT)
    r = _make("adaptive.yaml", case)
    oracle = make_oracle(case)
    target, max_inc, dt, interval, stop = 0.3, 2.0, 0.0002, 0.02, 0.12
    t, courant, dts = 0.0, None, []
    while not r.RDyFinished():
        r.RDyAdvance()
        # the oracle under the same rule
        if courant:
            dt = dt * min(target / courant, max_inc) if courant < target else dt * target / courant
            dt = min(dt, interval) if courant < target else dt
        dts.append(dt)
        t_end = t + interval
        nfull = int(math.floor((t_end - t) / dt * (1 + 1e-14) + 1e-12))
        rest = (t_end - t) - nfull * dt
        if rest < 1e-12 * max(t_end - t, dt):
            rest = 0.0
        for _ in range(nfull):
            oracle.rhs(dt)
            oracle.euler_update(dt)
        if rest:
            oracle.rhs(rest)
            oracle.euler_update(rest)
        courant = oracle.courant()
        t = t_end
        assert r.RDyGetTimeStep() == pytest.approx(dt, rel=1e-9)
        assert r.max_courant_number() == pytest.approx(courant, rel=1e-9)
    assert t == pytest.approx(stop) and len(dts) == 6 and dts[1] > dts[0]
    assert_state_close(r.state(), oracle.array("u_global"), what="adaptive RDyAdvance")
    r.RDyDestroy()


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [(0, 0), (3, 224)], ids=["K1", "K4"])
@pytest.mark.parametrize("deck", ["adaptive.yaml", "dam_break.yaml"])
def test_device_side_loop_needs_no_host_sync_and_equals_the_host_loop(deck, kernel, monkeypatch):
    """the adaptive-dt rule (src/rdyadvance.c:303-343), the step count of the interval and the shortened last step
    run in a device-side loop (one CUDA graph with a WHILE node per RDyAdvance; dt lives in device memory and the step kernels
    read it through a pointer).  A run of RDyAdvance calls therefore performs NO host synchronisation at all, and gives the same
    bits, the same dt and the same step count as the host-driven loop (RDYB_NO_DEVICE_LOOP=1), which reads the Courant number
    back once per interval."""
    from rdycore_b200.operator import host_sync_count

    case = cases.floodplain(nx=72, ny=56, eta0=40.0)  # fully wet (see test_adaptive_time_step_follows_the_reference_rule)
    rng = np.random.default_rng(5)
    case.u0[:, 1] = 0.4 * rng.normal(size=case.u0.shape[0]) * case.u0[:, 0]

    def run(host_loop):
        if host_loop:
            monkeypatch.setenv("RDYB_NO_DEVICE_LOOP", "1")
        else:
            monkeypatch.delenv("RDYB_NO_DEVICE_LOOP", raising=False)
        r = rdy.RDy(os.path.join(DECKS, deck))
        arr, nb = case.boundary_views()
        r.set_mesh(case.mesh, arr, nb)
        if kernel[0]:
            r.set_kernel(*kernel)
        r.RDySetup()
        r.RDySetDomainManningsN(case.mannings)
        r.RDySetDomainWaterSource(case.sources[:, 0])
        r.RDySetInitialConditions(case.u0)
        r.RDyAdvance()  # the first interval builds the graph (and may wait while doing so)
        n0, calls = host_sync_count(), 1
        while not r.RDyFinished():
            r.RDyAdvance()
            calls += 1
        syncs = host_sync_count() - n0
        out = (r.state(), r.RDyGetTimeStep(), r.RDyGetStep(), r.RDyGetTime(), r.max_courant_number())
        r.RDyDestroy()
        return out, syncs, calls

    dev, dev_syncs, calls = run(False)
    host, host_syncs, _ = run(True)
    assert calls >= 4
    assert dev_syncs == 0, "RDyAdvance on the device-side loop waited for the device"
    if deck == "adaptive.yaml":
        assert host_syncs >= calls - 1  # the host loop reads the Courant number back every interval
    assert np.array_equal(dev[0], host[0])
    assert dev[1:4] == host[1:4]
    assert dev[4] == host[4]
    assert not np.isnan(dev[0]).any() and dev[2] > calls  # several steps per interval


@pytest.mark.gpu
def test_setters_between_advances():
    """E3SM-style coupling: sources, Manning's n and Dirichlet values change between RDyAdvance calls
    (driver/main.c time loop; src/rdydata.c setters)."""
    from rdycore_b200.mesh import HostMesh

    mesh = HostMesh.quads(30, 12, 15.0, 6.0)
    bnd = cases.boundaries_from_predicates(mesh, [
        (1, "west", cases.DIRICHLET, lambda x, y: x < 1e-9),
        (2, "east", cases.CRITICAL_OUTFLOW, lambda x, y: x > 15.0 - 1e-9),
    ])
    n = mesh.num_owned_cells
    u0 = np.zeros((n, 3))
    u0[:, 0] = 1.0
    ne = len(bnd[0].edge_ids)
    dvals = np.tile(np.array([1.2, 0.5, 0.0]), (ne, 1))
    case = cases.Case("channel", mesh, cases.make_config(source_method=cases.IMPLICIT_XQ2018), bnd, u0, np.full(n, 0.03), np.zeros((n, 3)), 1e-3,
                      dirichlet={0: dvals})
    left = np.arange(n)[mesh.array("cell_xc")[:n] < 7.5].astype(np.int32)
    r = _make("dam_break.yaml", case, regions=[left])
    r.RDySetFlowDirichletBoundaryValues(0, ne, 3, dvals)
    oracle = make_oracle(case)
    rng = np.random.default_rng(5)
    for k in range(4):
        rain = 1e-3 * rng.random(len(left))
        nman = 0.02 + 0.02 * rng.random(n)
        d = dvals * (1 + 0.1 * k)
        r.RDySetRegionalWaterSource(0, rain)
        r.RDySetDomainManningsN(nman)
        r.RDySetFlowDirichletBoundaryValues(0, ne, 3, d)
        src = oracle.array("external_sources").reshape(-1, 3)
        src[left, 0] = rain
        oracle.array("material_properties")[:] = nman
        oracle.array("boundary_values", 0)[:] = d.ravel()
        r.RDyAdvance()
        oracle.step_euler(1e-3, 50)
    assert r.RDyFinished()
    assert_state_close(r.state(), oracle.array("u_global"), what="coupled RDyAdvance")
    # argument checks of the setters (src/rdydata.c:88-106, 225-240)
    with pytest.raises(RdybError) as ei:
        r.RDySetFlowDirichletBoundaryValues(1, len(bnd[1].edge_ids), 3, np.zeros((len(bnd[1].edge_ids), 3)))
    assert "but it has a different type" in str(ei.value)
    with pytest.raises(RdybError) as ei:
        r.RDySetFlowDirichletBoundaryValues(0, ne, 2, np.zeros((ne, 2)))
    assert "need to be three" in str(ei.value)
    with pytest.raises(RdybError) as ei:
        r.RDySetRegionalWaterSource(0, np.zeros(3))
    assert "Wrong size (3) for owned cells in region 0" in str(ei.value)
    with pytest.raises(RdybError):
        r.RDySetRegionalWaterSource(3, rain)
    r.RDyDestroy()


@pytest.mark.gpu
def test_restart_from_a_checkpoint_continues_bitwise(tmp_path):
    """The reference's restart test: run, checkpoint, restart from the file in a fresh RDy, finish — identical to the uninterrupted run
    (src/checkpoint.c:153-212, 258-320; GPU repeat runs are bitwise identical, so the comparison is exact)."""
    from rdycore_b200 import io

    case = cases.dam_break_quads(nx=50, ny=50, dt=1e-3)
    straight = _make("dam_break.yaml", case)
    while not straight.RDyFinished():
        straight.RDyAdvance()
    first = _make("dam_break.yaml", case)
    first.RDyAdvance()
    first.RDyAdvance()
    base = io.checkpoint_filename_base(str(tmp_path), "dam_break", first.RDyGetStep(), 200)
    assert base.endswith("dam_break.rdycore.r.100")
    path = first.RDyB200WriteCheckpoint(base)
    u_mid = first.state()
    first.RDyDestroy()
    ck = io.read_checkpoint(path)
    assert (ck["nproc"], ck["step"]) == (1, 100) and ck["t"] == pytest.approx(0.1) and ck["dt"] == pytest.approx(1e-3)
    assert np.array_equal(ck["solution"], u_mid)
    second = _make("dam_break.yaml", case)
    second.RDyB200ReadCheckpoint(path)
    assert second.RDyGetStep() == 100 and second.RDyGetTime() == pytest.approx(0.1)
    n_adv = 0
    while not second.RDyFinished():
        second.RDyAdvance()
        n_adv += 1
    assert n_adv == 2 and second.RDyGetStep() == 200
    assert np.array_equal(second.state(), straight.state())
    # restart.reinitialize: the state of the file, the clock back at zero
    third = _make("dam_break.yaml", case)
    third.RDyB200ReadCheckpoint(path, reinitialize=True)
    assert third.RDyGetStep() == 0 and third.RDyGetTime() == 0.0 and np.array_equal(third.state(), u_mid)
    with pytest.raises(RdybError, match="Invalid checkpoint file"):
        third.RDyB200ReadCheckpoint(str(tmp_path / "nope.h5"))
    for r in (straight, second, third):
        r.RDyDestroy()
