"""Forcing and boundary datasets (include/rdyb200_forcing.h) on the reference's own forcing fixtures:
Houston1km homogeneous rain / boundary series, two hourly Harvey rain rasters (77 x 38 pixels of 1 km), the Simons unstructured
rain and boundary datasets (334 and 377 points) — tests/golden/ref_fixtures/, copied by tests/golden/make_ref_fixtures.py.

CPU: the restated oracle == the reference's own rdyforcing_dataset.c / rdyforcing_map.c (compiled from /root/reference into
oracle/_ref), bit for bit; the host routine rdyb_forcing_current_data == oracle.
GPU: device index maps == oracle maps exactly (integers); values written into the operator's source / Dirichlet arrays == the
oracle's host arrays exactly (verbatim doubles, one multiplication); a run forced through the device path == the same run
forced through the host setters, bitwise."""
import os

import numpy as np
import pytest

from rdycore_b200 import cases, forcing, io
from rdycore_b200.operator import Operator

from oracle import forcing_binding
from test_reference_decks import deck_case

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "ref_fixtures")


def _vec(name):
    return io.read_petsc_vec(os.path.join(FIX, name))


def _houston():
    case = deck_case("Houston1km.DirichletBC")
    n = case.mesh.num_owned_cells
    return case, case.mesh.array("cell_xc")[:n].copy(), case.mesh.array("cell_yc")[:n].copy()


def _boundary_centroids(case, b):
    e = case.boundaries[b].edge_ids
    return case.mesh.array("edge_xc")[e].copy(), case.mesh.array("edge_yc")[e].copy()


def _scattered_targets(n, x0, x1, y0, y1, seed):
    rng = np.random.default_rng(seed)
    return rng.uniform(x0, x1, n), rng.uniform(y0, y1, n)


# ------------------------------------------------------------------ CPU: oracle pin + host routine
def _both():
    if not forcing_binding.available("reference"):
        pytest.skip("oracle/_ref/libforcing_oracle_ref.so not built (needs /root/reference at build time)")
    return forcing_binding.ForcingOracle("reference"), forcing_binding.ForcingOracle("port")


def test_port_equals_reference_forcing_code():
    ref, port = _both()
    rain, bc = _vec("Houston1km.rain.int32.bin"), _vec("Houston1km.bc.int32.bin")
    times = np.concatenate([np.linspace(-10.0, 5.2e5, 301), rain[0:40:2], bc[0:40:2]])
    for pairs in (rain, bc):
        for interp in (False, True):
            for t in times:
                assert ref.current_data(pairs, t, interp) == port.current_data(pairs, t, interp)
    # homogeneous datasets: the fill happens only when the bracket changes (or always when interpolating)
    for boundary in (0, 1):
        for interp in (False, True):
            a, b = np.full(21, -1.0), np.full(21, -1.0)
            sa, sb = [-1, -1], [-1, -1]
            for t in (0.0, 100.0, 950.0, 1800.0, 1e7):
                ref.homogeneous(bc, t, interp, boundary, a, sa)
                port.homogeneous(bc, t, interp, boundary, b, sb)
                assert np.array_equal(a, b) and sa == sb
    # maps: raster over the Houston cells and over scattered points (some far outside the raster), unstructured Simons data
    case, xc, yc = _houston()
    r = _vec("2017-08-26:00-00.int32.bin")
    ncols, nrows, xlc, ylc, cs = int(r[0]), int(r[1]), r[2], r[3], r[4]
    sx, sy = _scattered_targets(4000, xlc - 30 * cs, xlc + (ncols + 30) * cs, ylc - 30 * cs, ylc + (nrows + 30) * cs, 5)
    for mx, my in ((xc, yc), (sx, sy)):
        mr, mp = ref.map_raster(mx, my, ncols, nrows, xlc, ylc, cs), port.map_raster(mx, my, ncols, nrows, xlc, ylc, cs)
        assert np.array_equal(mr, mp) and mr.max() < ncols * nrows
        assert np.array_equal(ref.set_raster(r, 5, mr), port.set_raster(r, 5, mp))
    assert len(np.unique(ref.map_raster(xc, yc, ncols, nrows, xlc, ylc, cs))) > 500  # the mesh really spans the raster
    for pts, dat, stride in (("simons_forcing_x_y.int32.bin", "simons-ss.2011-08-27:14-00.int32.bin", 1), ("simons_bc_x_y.int32.bin", "simons-bc.2011-08-27:14-00.int32.bin", 3)):
        p, d = _vec(pts), _vec(dat)
        nd = int(p[0])
        assert int(p[1]) == 2 and int(d[0]) == nd and int(d[1]) == stride
        dx, dy = p[2::2][:nd], p[3::2][:nd]
        mx, my = _scattered_targets(3000, dx.min() - 500, dx.max() + 500, dy.min() - 500, dy.max() + 500, 9)
        mr, mp = ref.map_unstructured(mx, my, dx, dy), port.map_unstructured(mx, my, dx, dy)
        assert np.array_equal(mr, mp) and len(np.unique(mr)) > 100
        assert np.array_equal(ref.set_unstructured(d, stride, mr), port.set_unstructured(d, stride, mp))


def test_host_current_data_equals_the_oracle():
    oracle = forcing_binding.ForcingOracle()
    rain = _vec("Houston1km.rain.int32.bin")
    assert rain.size == 2 * 139 and rain[2] == 3600.0
    for interp in (False, True):
        for t in np.concatenate([np.linspace(-1.0, 5.1e5, 257), rain[0::2], rain[0::2] - 1e-9]):
            assert forcing.current_data(rain, t, interp) == oracle.current_data(rain, t, interp)


# ------------------------------------------------------------------ GPU
@pytest.mark.gpu
def test_device_maps_equal_the_oracle_maps():
    oracle = forcing_binding.ForcingOracle()
    case, xc, yc = _houston()
    op = Operator.from_case(case)
    r = _vec("2017-08-26:00-00.int32.bin")
    ncols, nrows, xlc, ylc, cs = int(r[0]), int(r[1]), r[2], r[3], r[4]
    sx, sy = _scattered_targets(20000, xlc - 30 * cs, xlc + (ncols + 30) * cs, ylc - 30 * cs, ylc + (nrows + 30) * cs, 5)
    # points exactly on pixel borders and corners: ties, resolved by scan order
    gx, gy = np.meshgrid(xlc + cs * np.arange(0, ncols + 1, 3), ylc + cs * np.arange(0, nrows + 1, 2))
    for mx, my in ((xc, yc), (sx, sy), (gx.ravel(), gy.ravel()), (gx.ravel() + cs / 2, gy.ravel())):
        m = forcing.DatasetMap.raster(op, mx, my, ncols, nrows, xlc, ylc, cs)
        assert np.array_equal(m.indices(), oracle.map_raster(mx, my, ncols, nrows, xlc, ylc, cs))
    for pts in ("simons_forcing_x_y.int32.bin", "simons_bc_x_y.int32.bin"):
        p = _vec(pts)
        nd = int(p[0])
        dx, dy = p[2::2][:nd], p[3::2][:nd]
        mx, my = _scattered_targets(20000, dx.min() - 500, dx.max() + 500, dy.min() - 500, dy.max() + 500, 9)
        mx[:nd], my[:nd] = dx, dy  # targets on top of data points
        m = forcing.DatasetMap.unstructured(op, mx, my, dx, dy)
        want = oracle.map_unstructured(mx, my, dx, dy)
        assert np.array_equal(m.indices(), want) and np.array_equal(want[:nd], np.arange(nd))


@pytest.mark.gpu
def test_raster_rain_and_unstructured_data_reach_the_operator_arrays():
    oracle = forcing_binding.ForcingOracle()
    case, xc, yc = _houston()
    op = Operator.from_case(case)
    n = op.num_owned
    for hour in ("00", "01"):  # RDyForcingSetRasterData, then the next hourly file
        r = _vec(f"2017-08-26:{hour}-00.int32.bin")
        ncols, nrows, xlc, ylc, cs = int(r[0]), int(r[1]), r[2], r[3], r[4]
        m = forcing.DatasetMap.raster(op, xc, yc, ncols, nrows, xlc, ylc, cs)
        forcing.set_source_from_dataset(op, 0, m, r, offset=5, stride=1, scale=forcing.MM_PER_HR_TO_M_PER_S)
        want = oracle.set_raster(r, 5, oracle.map_raster(xc, yc, ncols, nrows, xlc, ylc, cs))
        got = op.get_external_source(0)
        assert np.array_equal(got, want) and got.max() > 1e-7 and len(np.unique(got)) > 300
    # a map read from a file (-raster_rain_read_map_file) gives the same
    m2 = forcing.DatasetMap.from_indices(op, m.indices(), ncols * nrows)
    forcing.fill_source(op, 0, -1.0)
    assert np.array_equal(op.get_external_source(0), np.full(n, -1.0))  # homogeneous / constant datasets
    forcing.set_source_from_dataset(op, 0, m2, r, offset=5, stride=1, scale=forcing.MM_PER_HR_TO_M_PER_S)
    assert np.array_equal(op.get_external_source(0), want)
    # unstructured rain (stride 1) on the cells; unstructured boundary data (stride 3) on the Dirichlet boundary's edges
    p, d = _vec("simons_forcing_x_y.int32.bin"), _vec("simons-ss.2011-08-27:14-00.int32.bin")
    nd = int(p[0])
    # the Simons points lie elsewhere on the globe: move them over the Houston mesh, keeping their pattern
    dx = xc.min() + (p[2::2][:nd] - p[2::2][:nd].min()) / np.ptp(p[2::2][:nd]) * np.ptp(xc)
    dy = yc.min() + (p[3::2][:nd] - p[3::2][:nd].min()) / np.ptp(p[3::2][:nd]) * np.ptp(yc)
    mu = forcing.DatasetMap.unstructured(op, xc, yc, dx, dy)
    forcing.set_source_from_dataset(op, 0, mu, d, offset=2, stride=1, scale=1.0)
    want_u = oracle.set_unstructured(d, 1, oracle.map_unstructured(xc, yc, dx, dy))
    assert np.array_equal(op.get_external_source(0), want_u) and len(np.unique(want_u)) > 50
    b = next(i for i, bd in enumerate(case.boundaries) if bd.bc_type == cases.DIRICHLET)
    bx, by = _boundary_centroids(case, b)
    p, d = _vec("simons_bc_x_y.int32.bin"), _vec("simons-bc.2011-08-27:14-00.int32.bin")
    nd = int(p[0])
    dx = xc.min() + (p[2::2][:nd] - p[2::2][:nd].min()) / np.ptp(p[2::2][:nd]) * np.ptp(xc)
    dy = yc.min() + (p[3::2][:nd] - p[3::2][:nd].min()) / np.ptp(p[3::2][:nd]) * np.ptp(yc)
    mb = forcing.DatasetMap.unstructured(op, bx, by, dx, dy)
    forcing.set_boundary_values_from_dataset(op, b, mb, d, offset=2)
    want_b = oracle.set_unstructured(d, 3, oracle.map_unstructured(bx, by, dx, dy)).reshape(-1, 3)
    assert np.array_equal(forcing.get_boundary_values(op, b, len(bx)), want_b)
    forcing.fill_boundary_values(op, b, 1.25)  # RDyForcingSetHomogeneousBoundary: {h, 0, 0}
    assert np.array_equal(forcing.get_boundary_values(op, b, len(bx)), np.tile([1.25, 0.0, 0.0], (len(bx), 1)))
    with pytest.raises(RuntimeError, match="too short"):
        forcing.set_source_from_dataset(op, 0, m, r[:100], offset=5)
    with pytest.raises(RuntimeError, match="covers"):
        forcing.set_source_from_dataset(op, 0, mb, r, offset=5)


@pytest.mark.gpu
def test_houston_run_forced_on_the_device_equals_the_host_forced_run():
    """houston_hetero + homogeneous_bc of driver/tests/swe_roe/CMakeLists.txt: hourly rain raster, boundary depth series interpolated
    in time, 60 s coupling interval, dt 30 s.  One operator is forced the reference's way (host arrays through the setters, values
    from the oracle), the other through the device datasets: same states, bitwise."""
    oracle = forcing_binding.ForcingOracle()
    case, xc, yc = _houston()
    b = next(i for i, bd in enumerate(case.boundaries) if bd.bc_type == cases.DIRICHLET)
    ne = len(case.boundaries[b].edge_ids)
    bc = _vec("Houston1km.bc.int32.bin")
    host, dev = Operator.from_case(case), Operator.from_case(case)
    rasters = {h: _vec(f"2017-08-26:{h:02d}-00.int32.bin") for h in (0, 1)}
    r0 = rasters[0]
    ncols, nrows, xlc, ylc, cs = int(r0[0]), int(r0[1]), r0[2], r0[3], r0[4]
    idx = oracle.map_raster(xc, yc, ncols, nrows, xlc, ylc, cs)
    dmap = forcing.DatasetMap.raster(dev, xc, yc, ncols, nrows, xlc, ylc, cs)
    bvals, state = np.zeros(3 * ne), [-1, -1]
    t = 0.0
    for interval in range(80):  # 4800 s: crosses into the second hourly raster
        r = rasters[min(int(t // 3600.0), 1)]
        host.set_external_source(0, oracle.set_raster(r, 5, idx))
        oracle.homogeneous(bc, t, True, 1, bvals, state)
        host.set_boundary_values(b, bvals.reshape(-1, 3))
        forcing.set_source_from_dataset(dev, 0, dmap, r, offset=5, stride=1, scale=forcing.MM_PER_HR_TO_M_PER_S)
        forcing.fill_boundary_values(dev, b, forcing.current_data(bc, t, True)[1])
        host.step_euler(case.dt, 2)
        dev.step_euler(case.dt, 2)
        t += 2 * case.dt
    uh, ud = host.get_state(), dev.get_state()
    assert not np.isnan(uh).any() and np.array_equal(uh, ud)
    assert np.abs(uh - case.u0).max() > 1e-3


@pytest.mark.gpu
def test_observation_sites_and_boundary_flux_reset():
    """time_series.c: observations are the state at a few cells; accumulated boundary fluxes restart from zero after each write."""
    case, _, _ = _houston()
    b = next(i for i, bd in enumerate(case.boundaries) if bd.bc_type == cases.DIRICHLET)
    ne = len(case.boundaries[b].edge_ids)
    op = Operator.from_case(case)
    op.set_options(courant=True, boundary_fluxes=True)
    op.step_euler(case.dt, 20)
    u = op.get_state()
    sites = np.array([0, 17, 1234, op.num_owned - 1, 17], np.int32)
    assert np.array_equal(forcing.get_state_at(op, sites), u[sites])
    with pytest.raises(RuntimeError, match="out of range"):
        forcing.get_state_at(op, [op.num_owned])
    acc = op.get_boundary_fluxes(b, ne, accumulated=True)
    assert np.abs(acc).max() > 0.0
    forcing.reset_boundary_fluxes(op)
    assert np.abs(op.get_boundary_fluxes(b, ne, accumulated=True)).max() == 0.0
    op.step_euler(case.dt, 1)
    one = op.get_boundary_fluxes(b, ne, accumulated=True)
    inst = op.get_boundary_fluxes(b, ne, accumulated=False)
    assert np.allclose(one, case.dt * inst, rtol=1e-14, atol=0.0)
