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
                ("total_edges", C.c_int64), ("total_halo", C.c_int64), ("edge_off", c_f64p)]


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


def _check_offsets(case, tile_cells, halo=None):
    view, L, T = _layouts(case, tile_cells, halo)
    try:
        va = case.mesh.view_arrays()
        xc, yc, xm, ym, ecells = va["cell_xc"], va["cell_yc"], va["edge_xm"], va["edge_ym"], va["edge_cell_ids"].reshape(-1, 2)
        plane = T.total_edges
        off = _np(T.edge_off, 4 * plane).reshape(4, plane)
        desc = _np(T.tile_desc, 8 * T.num_tiles).reshape(-1, 8)
        lr, eid, bidx = _np(T.edge_lr, plane), _np(T.edge_id, plane), _np(T.edge_bidx, plane)
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


def test_tile_edge_face_offsets_with_ghosts():
    """A 2-rank block partition: frontier tiles are short and list ghost cells in their halo; offsets of owned-ghost edges included."""
    case = cases.floodplain(nx=40, ny=24, eta0=40.0, px=2, py=1, rank=1)
    case.config.reconstruction = cases.MUSCL_VANLEER
    halo = case.mesh.halo()
    n_int, n_bnd, nt = _check_offsets(case, 128, halo)
    assert n_int > 0 and n_bnd > 0 and nt > 4


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
