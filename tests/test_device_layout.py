"""CPU: the tiled device layout and its second-order face offsets (rdycore_b200/csrc/device_layout.c), checked against the mesh arrays.

The tile edges carry what ReconstructFaceValues (src/operator_fluxes_ceed.c:1162-1172) reads per internal edge: edge midpoint minus
the centroid of the left and of the right cell.  No compute call: only the host-side layout builders run."""
import ctypes as C

import numpy as np
import pytest

from rdycore_b200 import cases
from rdycore_b200._lib import c_f64p, c_i32p, check, lib

c_u16p, c_u32p = C.POINTER(C.c_uint16), C.POINTER(C.c_uint32)


class RdybLayout(C.Structure):  # device_layout.h
    _fields_ = [("n_owned", C.c_int32), ("n_local", C.c_int32), ("n_interior", C.c_int32), ("stride", C.c_int32), ("width", C.c_int32),
                ("num_view_cells", C.c_int32), ("dev_to_local", c_i32p), ("local_to_dev", c_i32p), ("owned_to_dev", c_i32p), ("dev_to_owned", c_i32p),
                ("inv_area", c_f64p), ("area", c_f64p), ("dzdx", c_f64p), ("dzdy", c_f64p), ("zc", c_f64p),
                ("ell_nbr", c_i32p), ("ell_sn", c_f64p), ("ell_cn", c_f64p), ("ell_slen", c_f64p), ("ell_edge", c_i32p),
                ("ell_gcx", c_f64p), ("ell_gcy", c_f64p), ("ell_dxs", c_f64p), ("ell_dys", c_f64p), ("ell_dxn", c_f64p), ("ell_dyn", c_f64p),
                ("num_boundaries", C.c_int32), ("num_boundary_edges", C.c_int32), ("b_offsets", c_i32p), ("b_type", c_i32p), ("b_cell", c_i32p),
                ("num_peers", C.c_int32), ("peer_ranks", c_i32p), ("send_offsets", c_i32p), ("send_cells", c_i32p), ("recv_offsets", c_i32p)]


class RdybTiledLayout(C.Structure):
    _fields_ = [("tile_cells", C.c_int32), ("num_tiles", C.c_int32), ("num_interior_tiles", C.c_int32), ("max_halo", C.c_int32), ("max_edges", C.c_int32),
                ("max_halo_interior", C.c_int32), ("max_edges_interior", C.c_int32),
                ("tile_cell_begin", c_i32p), ("tile_desc", c_i32p), ("halo_offsets", c_i32p), ("halo_cells", c_i32p), ("edge_offsets", c_i32p),
                ("edge_lr", c_u32p), ("edge_bidx", c_i32p), ("edge_id", c_i32p), ("edge_cell", c_i32p),
                ("edge_sn", c_f64p), ("edge_cn", c_f64p), ("edge_len", c_f64p), ("cell_edges", c_u16p),
                ("total_edges", C.c_int64), ("total_halo", C.c_int64), ("edge_off", c_f64p), ("edge_off_plane", C.c_int64),
                ("halo2_cells", c_i32p), ("halo_nbr", c_u16p), ("max_halo2", C.c_int32), ("two_ring", C.c_int32), ("total_halo2", C.c_int64)]


def _np(ptr, n):
    return np.ctypeslib.as_array(ptr, shape=(n,))


def _layouts(case, tile_cells, halo=None):
    view = case.mesh.view()
    arr, nb = case.boundary_views()
    L, T = RdybLayout(), RdybTiledLayout()
    check(lib().rdyb_layout_build(C.byref(view), C.c_int32(nb), arr, C.byref(halo) if halo is not None else None, C.c_int(1), C.byref(L)))
    check(lib().rdyb_layout_build_muscl(C.byref(view), C.byref(L)))
    check(lib().rdyb_tiled_layout_build(C.byref(L), C.c_int32(tile_cells), C.byref(T)))
    check(lib().rdyb_tiled_layout_build_muscl(C.byref(L), C.byref(T)))
    return view, L, T


def _layouts_first_order(case, tile_cells, halo=None):
    view = case.mesh.view()
    arr, nb = case.boundary_views()
    L, T = RdybLayout(), RdybTiledLayout()
    check(lib().rdyb_layout_build(C.byref(view), C.c_int32(nb), arr, C.byref(halo) if halo is not None else None, C.c_int(1), C.byref(L)))
    check(lib().rdyb_tiled_layout_build(C.byref(L), C.c_int32(tile_cells), C.byref(T)))
    return view, L, T


def _check_offsets(case, tile_cells, halo=None):
    view, L, T = _layouts(case, tile_cells, halo)
    try:
        va = case.mesh.view_arrays()
        xc, yc, xm, ym, ecells = va["cell_xc"], va["cell_yc"], va["edge_xm"], va["edge_ym"], va["edge_cell_ids"].reshape(-1, 2)
        plane, ne = T.edge_off_plane, T.total_edges
        assert plane % 4 == 0 and 0 <= plane - ne < 4  # every plane starts 16-byte aligned (bulk copies of a tile's slice)
        off = _np(T.edge_off, 4 * plane).reshape(4, plane)
        desc = _np(T.tile_desc, 8 * T.num_tiles).reshape(-1, 8)
        lr, eid, bidx = _np(T.edge_lr, ne), _np(T.edge_id, ne), _np(T.edge_bidx, ne)
        halo_cells, d2l = _np(T.halo_cells, T.total_halo), _np(L.dev_to_local, L.n_local)
        n_int = n_bnd = 0
        for cb, ncell, eo, nedge, ho, nhalo, _, _ in desc:
            for g in range(eo, eo + nedge):
                l, r = int(lr[g]) & 0xFFFF, int(lr[g]) >> 16
                if r == 0xFFFF:
                    assert bidx[g] >= 0 and not off[:, g].any()  # boundary edges stay first order
                    n_bnd += 1
                    continue
                dev = [cb + x if x < ncell else halo_cells[ho + x - T.tile_cells] for x in (l, r)]
                loc = [d2l[d] for d in dev]
                e = eid[g]
                assert list(ecells[e]) == loc  # tile edge orientation = the mesh's left / right cells
                want = [xm[e] - xc[loc[0]], ym[e] - yc[loc[0]], xm[e] - xc[loc[1]], ym[e] - yc[loc[1]]]
                assert list(off[:, g]) == want  # bitwise: the same subtraction on the same doubles
                n_int += 1
        return n_int, n_bnd, T.num_tiles
    finally:
        lib().rdyb_tiled_layout_free(C.byref(T))
        lib().rdyb_layout_free(C.byref(L))


@pytest.mark.parametrize("tile_cells", [128, 224])
def test_tile_edge_face_offsets_triangles(tile_cells):
    case = cases.floodplain(nx=24, ny=20, eta0=40.0)
    case.config.reconstruction = cases.MUSCL_MINMOD
    n_int, n_bnd, nt = _check_offsets(case, tile_cells)
    n_cells = 2 * 24 * 20
    assert nt == -(-n_cells // tile_cells) and n_bnd == 2 * (24 + 20)
    assert n_int >= (3 * n_cells - n_bnd) // 2  # every internal edge once, edges between two tiles twice


def test_tile_edge_face_offsets_with_ghosts(monkeypatch):
    """A 2-rank block partition: frontier tiles are short and list ghost cells in their halo; offsets of owned-ghost edges included.
    Built with the opt-in edge schedule, which permutes the tile edges after they were created."""
    monkeypatch.setenv("RDYB_EDGE_SCHEDULE", "1")
    case = cases.floodplain(nx=40, ny=24, eta0=40.0, px=2, py=1, rank=1)
    case.config.reconstruction = cases.MUSCL_VANLEER
    halo = case.mesh.halo()
    n_int, n_bnd, nt = _check_offsets(case, 128, halo)
    assert n_int > 0 and n_bnd > 0 and nt > 4


@pytest.mark.parametrize("which,tile_cells", [("triangles", 192), ("triangles", 128), ("quads", 192)])
def test_two_ring_layout_invariants(which, tile_cells):
    """K4M's two-ring staging (rdyb_tiled_layout_build_two_ring): every interior neighbour of every halo cell of a tile is staged - as a
    tile cell, a halo cell or a second-ring cell - at the position halo_nbr names, in ELL slot order; the second ring is sorted, disjoint
    from tile and halo, and its slices are 16-byte aligned."""
    case = cases.floodplain(nx=44, ny=38, eta0=8.0) if which == "triangles" else cases.dam_break_quads(nx=40, ny=30)
    view, L, T = _layouts(case, tile_cells)
    try:
        assert T.two_ring == 1 and T.max_halo2 > 0
        S, W, TC = L.stride, L.width, T.tile_cells
        nbr = _np(L.ell_nbr, 4 * S).reshape(4, S)
        desc = _np(T.tile_desc, 8 * T.num_tiles).reshape(-1, 8)
        halo, halo2 = _np(T.halo_cells, T.total_halo), _np(T.halo2_cells, T.total_halo2)
        hn = _np(T.halo_nbr, 4 * T.total_halo).reshape(-1, 4)
        worst = 0
        for cb, ncell, _, _, ho, nh, h2o, nh2 in desc:
            assert h2o % 4 == 0 and 0 <= nh2 <= T.max_halo2
            H, H2 = halo[ho:ho + nh], halo2[h2o:h2o + nh2]
            assert np.all(np.diff(H2) > 0) and not np.intersect1d(H, H2).size and not np.any((H2 >= cb) & (H2 < cb + ncell))
            used = np.zeros(nh2, bool)
            for j in range(nh):
                for k in range(4):
                    v = int(hn[ho + j, k])
                    n = int(nbr[k, H[j]]) if k < W else -1
                    if n < 0:
                        assert v == 0xFFFF  # boundary or empty slot (RDYB_NO_EDGE is negative too)
                    elif v & 0x8000:
                        assert H2[v & 0x7FFF] == n
                        used[v & 0x7FFF] = True
                    elif v >= TC:
                        assert H[v - TC] == n
                    else:
                        assert cb + v == n and v < ncell
            assert used.all()  # nothing is staged in vain
            worst = max(worst, nh2)
        assert worst == T.max_halo2
    finally:
        lib().rdyb_tiled_layout_free(C.byref(T))
        lib().rdyb_layout_free(C.byref(L))


def test_two_ring_staging_reproduces_the_cell_centric_gradients():
    """K4M's phase 0 restated in numpy on the layout arrays: per tile, stage tile | ring 1 | ring 2 states, derive the gradient of every
    tile cell through cell_edges + edge_lr and of every halo cell through halo_nbr, in ELL slot order - and compare with the cell-centric
    sum of muscl_gradient_kernel (weights with the side folded in, times q_nbr - q_self).  The same products in the same order: equal bits,
    for every tile that stages a cell.  Also pins the folded weights against the reference's w * (q_right - q_left) form."""
    case = cases.floodplain(nx=44, ny=38, eta0=8.0)
    view, L, T = _layouts(case, 192)
    try:
        S, W, N, TC = L.stride, L.width, L.n_owned, T.tile_cells
        assert W == 3
        nbr = _np(L.ell_nbr, 4 * S).reshape(4, S)
        gcx, gcy, slen = _np(L.ell_gcx, 4 * S).reshape(4, S), _np(L.ell_gcy, 4 * S).reshape(4, S), _np(L.ell_slen, 4 * S).reshape(4, S)
        rng = np.random.default_rng(5)
        q = rng.normal(size=L.n_local)

        def cell_centric(c):
            gx = gy = 0.0
            for k in range(W):
                n = int(nbr[k, c])
                if n < 0:
                    continue
                d = q[n] - q[c]
                gx, gy = gcx[k, c] * d + gx, gcy[k, c] * d + gy
            return gx, gy

        # folded weight * (q_nbr - q_self) == unfolded weight * (q_right - q_left): the cell is the left cell iff slen < 0
        for c in range(0, N, 37):
            for k in range(W):
                n = int(nbr[k, c])
                if n >= 0:
                    left = slen[k, c] < 0.0
                    w = gcx[k, c] if left else -gcx[k, c]
                    assert w * ((q[n] - q[c]) if left else (q[c] - q[n])) == gcx[k, c] * (q[n] - q[c])
        desc = _np(T.tile_desc, 8 * T.num_tiles).reshape(-1, 8)
        halo, halo2 = _np(T.halo_cells, T.total_halo), _np(T.halo2_cells, T.total_halo2)
        hn = _np(T.halo_nbr, 4 * T.total_halo).reshape(-1, 4)
        lr, ce = _np(T.edge_lr, T.total_edges), _np(T.cell_edges, 4 * S).reshape(4, S)
        HM = 80
        for cb, ncell, eo, nedge, ho, nh, h2o, nh2 in desc:
            staged = np.full(TC + HM + 96, np.nan)
            staged[:ncell] = q[cb:cb + ncell]
            staged[TC:TC + nh] = q[halo[ho:ho + nh]]
            staged[TC + HM:TC + HM + nh2] = q[halo2[h2o:h2o + nh2]]
            for t in range(ncell):  # tile cells: neighbour = the other endpoint of the slot's tile edge
                gx = gy = 0.0
                for k in range(W):
                    v = int(ce[k, cb + t])
                    if v == 0xFFFF:
                        continue
                    e = int(lr[eo + (v & 0x7FFF)])
                    if e >> 16 == 0xFFFF:
                        continue
                    n = (e & 0xFFFF) if v & 0x8000 else (e >> 16)
                    d = staged[n] - staged[t]
                    gx, gy = gcx[k, cb + t] * d + gx, gcy[k, cb + t] * d + gy
                assert (gx, gy) == cell_centric(cb + t)
            for j in range(nh):  # ring-1 cells: neighbours at the staged positions halo_nbr names
                c = int(halo[ho + j])
                gx = gy = 0.0
                for k in range(W):
                    v = int(hn[ho + j, k])
                    if v == 0xFFFF:
                        continue
                    n = TC + HM + (v & 0x7FFF) if v & 0x8000 else v
                    d = staged[n] - staged[TC + j]
                    gx, gy = gcx[k, c] * d + gx, gcy[k, c] * d + gy
                assert not np.isnan(gx) and (gx, gy) == cell_centric(c)
    finally:
        lib().rdyb_tiled_layout_free(C.byref(T))
        lib().rdyb_layout_free(C.byref(L))


def test_two_ring_layout_is_off_when_a_halo_cell_is_a_ghost():
    """On a partitioned mesh the first ring of a frontier tile holds ghosts, whose own neighbours are not all on this rank: no two-ring
    staging (the gradients of ghosts arrive by halo exchange, K4M runs with the gradient array)."""
    case = cases.floodplain(nx=44, ny=30, eta0=8.0, px=2, py=1, rank=1)
    halo = case.mesh.halo()
    view, L, T = _layouts(case, 192, halo)
    try:
        assert T.two_ring == 0 and not T.halo_nbr and not T.halo2_cells
    finally:
        lib().rdyb_tiled_layout_free(C.byref(T))
        lib().rdyb_layout_free(C.byref(L))


def test_tiled_muscl_layout_needs_the_slot_geometry():
    case = cases.floodplain(nx=16, ny=16, eta0=40.0)
    view = case.mesh.view()
    arr, nb = case.boundary_views()
    L, T = RdybLayout(), RdybTiledLayout()
    check(lib().rdyb_layout_build(C.byref(view), C.c_int32(nb), arr, None, C.c_int(1), C.byref(L)))
    check(lib().rdyb_tiled_layout_build(C.byref(L), C.c_int32(128), C.byref(T)))
    assert lib().rdyb_tiled_layout_build_muscl(C.byref(L), C.byref(T)) != 0
    assert b"slot geometry" in lib().rdyb_last_error()
    lib().rdyb_tiled_layout_free(C.byref(T))
    lib().rdyb_layout_free(C.byref(L))


NO_EDGE = -2**31


@pytest.mark.parametrize("schedule", [False, True])
@pytest.mark.parametrize("partitioned", [False, True])
@pytest.mark.parametrize("tile_cells", [128, 224])
def test_tiled_layout_invariants(tile_cells, partitioned, schedule, monkeypatch):
    """What the tiled kernels (K2-K4, K2M, K3M) assume about a tile: padded slices, a sorted halo outside the tile, edges with at least one
    tile cell, per-cell edge slots that mirror the ELL slots (same edge, same side, same geometry, same order), short frontier tiles."""
    if schedule:  # the opt-in bank-aware edge order permutes a tile's edges; every invariant must survive it
        monkeypatch.setenv("RDYB_EDGE_SCHEDULE", "1")
    if partitioned:
        case = cases.floodplain(nx=44, ny=30, eta0=8.0, px=2, py=1, rank=0)
        halo = case.mesh.halo()
    else:
        case, halo = cases.dam_break_quads(nx=30, ny=22), None
    view, L, T = _layouts_first_order(case, tile_cells, halo)
    try:
        S, W, TC, N, NI = L.stride, L.width, T.tile_cells, L.n_owned, L.n_interior
        plane = T.total_edges
        desc = _np(T.tile_desc, 8 * T.num_tiles).reshape(-1, 8)
        lr, eid, bidx = _np(T.edge_lr, plane), _np(T.edge_id, plane), _np(T.edge_bidx, plane)
        esn, ecn, elen = _np(T.edge_sn, plane), _np(T.edge_cn, plane), _np(T.edge_len, plane)
        halo_cells = _np(T.halo_cells, T.total_halo)
        cell_edges = _np(T.cell_edges, 4 * S).reshape(4, S)
        nbr, slen = _np(L.ell_nbr, 4 * S).reshape(4, S), _np(L.ell_slen, 4 * S).reshape(4, S)
        lsn, lcn, ledge = _np(L.ell_sn, 4 * S).reshape(4, S), _np(L.ell_cn, 4 * S).reshape(4, S), _np(L.ell_edge, 4 * S).reshape(4, S)
        assert NI < N if partitioned else NI == N
        covered = 0
        for t, (cb, ncell, eo, nedge, ho, nhalo, _, _) in enumerate(desc):
            interior = t < T.num_interior_tiles
            assert eo % 4 == 0 and ho % 4 == 0 and 0 < ncell <= TC and cb == covered
            covered += ncell
            assert (cb + ncell <= NI) if interior else (cb >= NI)
            # the producer warps fetch every per-tile slice with bulk copies rounded up to 16 bytes (K3, K4, K3M): the rounded
            # slices must stay inside the arrays, and start 16-byte aligned for every element size in use (2, 4, 8 bytes)
            assert cb % 8 == 0 and cb + ((ncell + 7) & ~7) <= S
            assert eo + ((nedge + 3) & ~3) <= T.total_edges and ho + nhalo <= T.total_halo
            h = halo_cells[ho:ho + nhalo]
            assert np.all(np.diff(h) > 0) and not np.any((h >= cb) & (h < cb + ncell)) and np.all((h >= 0) & (h < L.n_local))
            if interior:
                assert np.all(h < N)  # interior cells have no ghost neighbour
            else:
                assert ncell <= 80 // W and nhalo <= 80 and nedge <= 80  # frontier tiles fit every compiled tile shape
            used = np.zeros(nedge, bool)
            for i in range(cb, cb + ncell):
                for k in range(W):
                    ce = int(cell_edges[k, i])
                    if nbr[k, i] == NO_EDGE:
                        assert ce == 0xFFFF
                        continue
                    e, right = ce & 0x7FFF, bool(ce & 0x8000)
                    assert e < nedge and right == (slen[k, i] > 0)
                    g = eo + e
                    used[e] = True
                    l, r = int(lr[g]) & 0xFFFF, int(lr[g]) >> 16
                    assert (r if right else l) == i - cb
                    other = l if right else r
                    if nbr[k, i] < 0:  # boundary slot
                        assert other == 0xFFFF and not right and bidx[g] == -(nbr[k, i] + 1)
                    else:
                        dev = cb + other if other < ncell else h[other - TC]
                        assert dev == nbr[k, i] and bidx[g] == -1
                    assert eid[g] == ledge[k, i] and esn[g] == lsn[k, i] and ecn[g] == lcn[k, i] and elen[g] == abs(slen[k, i])
            assert used.all()  # no tile edge without a tile cell
        assert covered == N
    finally:
        lib().rdyb_tiled_layout_free(C.byref(T))
        lib().rdyb_layout_free(C.byref(L))


@pytest.mark.parametrize("which", ["quads", "triangles_rank0", "triangles_rank1", "mixed"])
def test_cell_centric_layout_mirrors_the_mesh(which):
    """The ELL layout every kernel is derived from (rdyb_layout_build): renumbering (interior, frontier, ghosts grouped by peer), and per
    slot the mesh edge with its orientation, normal, length — in the order in which the reference accumulates into a cell: interior
    edges by position in internal_edge_ids (src/swe/swe_petsc.c:244-310), then boundary edges by boundary (operator_fluxes_petsc.c:28-51)."""
    halo = None
    if which == "quads":
        case = cases.dam_break_quads(nx=26, ny=18)
    elif which == "mixed":
        ex = __import__("rdycore_b200.io", fromlist=["io"])
        import os
        e = ex.read_exodus(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_fixtures", "DamBreak_grid5x10_mixed_elements.exo"))
        mesh = ex.mesh_from_exodus(e)
        bnd = ex.boundaries_from_side_sets(mesh, e)
        n = mesh.num_owned_cells
        case = cases.Case("mixed", mesh, cases.make_config(), bnd, np.ones((n, 3)), np.full(n, 0.02), None, 1e-3)
    else:
        case = cases.floodplain(nx=30, ny=22, eta0=8.0, px=2, py=1, rank=int(which[-1]))
        halo = case.mesh.halo()
    view = case.mesh.view()
    arr, nb = case.boundary_views()
    L = RdybLayout()
    check(lib().rdyb_layout_build(C.byref(view), C.c_int32(nb), arr, C.byref(halo) if halo is not None else None, C.c_int(1), C.byref(L)))
    try:
        va = case.mesh.view_arrays()
        N, NI, NL, S, W = L.n_owned, L.n_interior, L.n_local, L.stride, L.width
        assert N == view.num_owned_cells and S % 32 == 0 and S >= NL and W == (3 if which.startswith("tri") else 4)
        d2l, l2d = _np(L.dev_to_local, NL), _np(L.local_to_dev, L.num_view_cells)
        assert len(set(d2l.tolist())) == NL and all(l2d[d2l[d]] == d for d in range(NL))
        owned = va["cell_is_owned"]
        assert owned[d2l[:N]].all() and not owned[d2l[N:]].any()
        o2d, d2o = _np(L.owned_to_dev, N), _np(L.dev_to_owned, N)
        assert np.array_equal(o2d[d2o], np.arange(N)) and np.array_equal(va["cell_local_to_owned"][d2l[:N]], d2o)
        nbr, slen = _np(L.ell_nbr, 4 * S).reshape(4, S), _np(L.ell_slen, 4 * S).reshape(4, S)
        sn, cn, eedge = _np(L.ell_sn, 4 * S).reshape(4, S), _np(L.ell_cn, 4 * S).reshape(4, S), _np(L.ell_edge, 4 * S).reshape(4, S)
        ecells, elen, esn, ecn = va["edge_cell_ids"].reshape(-1, 2), va["edge_lengths"], va["edge_sn"], va["edge_cn"]
        pos_in_internal = {int(e): p for p, e in enumerate(va["internal_edge_ids"])}
        assert np.array_equal(_np(L.inv_area, NL), 1.0 / va["cell_areas"][d2l]) and np.array_equal(_np(L.area, NL), va["cell_areas"][d2l])
        b_off, b_cell = _np(L.b_offsets, L.num_boundaries + 1), _np(L.b_cell, max(L.num_boundary_edges, 1))
        flat_edges = np.concatenate([np.asarray(b.edge_ids, dtype=np.int64) for b in case.boundaries]) if case.boundaries else np.zeros(0, np.int64)
        assert L.num_boundaries == nb and b_off[-1] == len(flat_edges) == L.num_boundary_edges
        seen_edges = extras = 0
        for i in range(N):
            me, last_int, last_bnd, ghost_nbrs = d2l[i], -1, -1, 0
            for k in range(4):
                if nbr[k, i] == NO_EDGE:
                    continue
                assert k < W
                e = eedge[k, i]
                seen_edges += 1
                assert abs(slen[k, i]) == elen[e] and sn[k, i] == esn[e] and cn[k, i] == ecn[e]
                if nbr[k, i] >= 0:
                    left = slen[k, i] < 0
                    assert list(ecells[e]) == ([me, d2l[nbr[k, i]]] if left else [d2l[nbr[k, i]], me])
                    assert last_bnd < 0 and pos_in_internal[int(e)] > last_int  # reference accumulation order
                    last_int = pos_in_internal[int(e)]
                    ghost_nbrs += nbr[k, i] >= N
                else:
                    j = -(nbr[k, i] + 1)
                    assert slen[k, i] < 0 and ecells[e][0] == me and ecells[e][1] < 0 and flat_edges[j] == e and b_cell[j] == i and j > last_bnd
                    last_bnd = j
            if i < NI:
                assert ghost_nbrs == 0  # interior cells run before the halo arrives: none may read a ghost
            else:
                extras += ghost_nbrs == 0
        # the interior range ends on a multiple of 16 cells (aligned slices for bulk copies): at most 15 interior cells ride with the frontier
        assert extras == 0 if NI == N else (NI % 16 == 0 and extras < 16)
        n_int_owned_sides = sum(int(owned[c]) for e in va["internal_edge_ids"] for c in ecells[e])
        n_bnd = sum(1 for j in range(len(flat_edges)) if b_cell[j] >= 0)
        assert seen_edges == n_int_owned_sides + n_bnd
        if halo is not None:  # ghosts sit after the owned cells, grouped by the rank that owns them, in the halo plan's receive order
            h = case.mesh.halo_arrays()
            ro = _np(L.recv_offsets, L.num_peers + 1)
            assert L.num_peers == len(h["peer_ranks"]) and np.array_equal(ro, h["recv_offsets"]) and N + ro[-1] == NL
            assert np.array_equal(d2l[N:], h["recv_cells"])
            assert np.array_equal(d2l[_np(L.send_cells, h["send_offsets"][-1])], h["send_cells"])
    finally:
        lib().rdyb_layout_free(C.byref(L))
