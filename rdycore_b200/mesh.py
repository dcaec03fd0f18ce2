"""Host mesh (C, rdyb200_mesh.h) behind a thin Python handle.

Mirrors what RDyMeshCreateFromDM (src/rdymesh.c:1401-1440) hands to the operator: all heavy
lifting is in csrc/host_mesh.c; this class only owns the handle and exposes numpy views.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import RdybHaloView, RdybMeshView, check, lib

_DTYPES = {
    "boundary_edge_ids": np.int32, "edge_vertex_ids": np.int32, "owned_to_local": np.int32, "cell_global_ids": np.int64,
    "cell_offsets": np.int32, "cell_vertex_ids": np.int32,
}


def _as(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype=dtype)
    ctype = {np.int32: C.c_int32, np.int64: C.c_int64, np.float64: C.c_double}[dtype]
    return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ctype)), shape=(n,))


class HostMesh:
    def __init__(self, handle: C.c_void_p):
        self._h = handle
        self._keep = []

    # ---- constructors ----
    @classmethod
    def from_cells(cls, xyz: np.ndarray, cell_offsets: np.ndarray, cell_verts: np.ndarray) -> "HostMesh":
        xyz = np.ascontiguousarray(xyz, dtype=np.float64).reshape(-1, 3)
        offs = np.ascontiguousarray(cell_offsets, dtype=np.int32)
        verts = np.ascontiguousarray(cell_verts, dtype=np.int32)
        h = C.c_void_p()
        check(lib().rdyb_hostmesh_create(C.c_int32(xyz.shape[0]), xyz.ctypes.data_as(C.c_void_p), C.c_int32(len(offs) - 1),
                                         offs.ctypes.data_as(C.c_void_p), verts.ctypes.data_as(C.c_void_p), C.byref(h)))
        return cls(h)

    @classmethod
    def quads(cls, nx: int, ny: int, lx: float, ly: float) -> "HostMesh":
        h = C.c_void_p()
        check(lib().rdyb_hostmesh_generate_quads(C.c_int32(nx), C.c_int32(ny), C.c_double(lx), C.c_double(ly), C.byref(h)))
        return cls(h)

    @classmethod
    def floodplain(cls, nx: int, ny: int, lx: float, ly: float, jitter: float = 0.2, zamp: float = 1.0, seed: int = 12345,
                   px: int = 1, py: int = 1, rank: int = 0) -> "HostMesh":
        h = C.c_void_p()
        check(lib().rdyb_hostmesh_generate_floodplain(C.c_int32(nx), C.c_int32(ny), C.c_double(lx), C.c_double(ly), C.c_double(jitter),
                                                      C.c_double(zamp), C.c_uint64(seed), C.c_int32(px), C.c_int32(py), C.c_int32(rank),
                                                      C.byref(h)))
        return cls(h)

    def refine_triangles(self) -> "HostMesh":
        h = C.c_void_p()
        check(lib().rdyb_hostmesh_refine_triangles(self._h, C.byref(h)))
        return HostMesh(h)

    def partition_rcb(self, nparts: int) -> np.ndarray:
        part = np.zeros(self.num_cells, dtype=np.int32)
        check(lib().rdyb_hostmesh_partition_rcb(self._h, C.c_int32(nparts), part.ctypes.data_as(C.c_void_p)))
        return part

    def extract_local(self, part: np.ndarray, nparts: int, rank: int) -> "HostMesh":
        part = np.ascontiguousarray(part, dtype=np.int32)
        h = C.c_void_p()
        check(lib().rdyb_hostmesh_extract_local(self._h, part.ctypes.data_as(C.c_void_p), C.c_int32(nparts), C.c_int32(rank), C.byref(h)))
        return HostMesh(h)

    # ---- modifiers ----
    def override_2d_projection(self) -> None:
        check(lib().rdyb_hostmesh_override_2d_projection(self._h))

    def set_zc(self, zc: np.ndarray | None = None) -> None:
        if zc is None:
            check(lib().rdyb_hostmesh_set_zc(self._h, None))
        else:
            zc = np.ascontiguousarray(zc, dtype=np.float64)
            assert zc.shape[0] == self.num_cells
            check(lib().rdyb_hostmesh_set_zc(self._h, zc.ctypes.data_as(C.c_void_p)))

    def set_vertex_z(self, z: np.ndarray) -> None:
        z = np.ascontiguousarray(z, dtype=np.float64)
        assert z.shape[0] == self.num_vertices
        check(lib().rdyb_hostmesh_set_vertex_z(self._h, z.ctypes.data_as(C.c_void_p)))

    # ---- accessors ----
    def sizes(self):
        out = (C.c_int32 * 6)()
        check(lib().rdyb_hostmesh_sizes(self._h, out))
        return list(out)

    num_vertices = property(lambda s: s.sizes()[0])
    num_cells = property(lambda s: s.sizes()[1])
    num_owned_cells = property(lambda s: s.sizes()[2])
    num_edges = property(lambda s: s.sizes()[3])
    num_internal_edges = property(lambda s: s.sizes()[4])
    num_boundary_edges = property(lambda s: s.sizes()[5])

    def array(self, name: str) -> np.ndarray:
        ptr, n = C.c_void_p(), C.c_int64()
        check(lib().rdyb_hostmesh_array(self._h, name.encode(), C.byref(ptr), C.byref(n)))
        return _as(ptr, n.value, _DTYPES.get(name, np.float64))

    def vertices(self) -> np.ndarray:
        ptr, n = C.c_void_p(), C.c_int32()
        check(lib().rdyb_hostmesh_vertices(self._h, C.byref(ptr), C.byref(n)))
        return _as(ptr, 3 * n.value, np.float64).reshape(-1, 3)

    def view(self) -> RdybMeshView:
        v = RdybMeshView()
        check(lib().rdyb_hostmesh_view(self._h, C.byref(v)))
        return v

    def halo(self) -> RdybHaloView:
        h = RdybHaloView()
        check(lib().rdyb_hostmesh_halo(self._h, C.byref(h)))
        return h

    def view_arrays(self) -> dict:
        """numpy views of every array in the RdybMeshView (borrowed; valid while the mesh lives)."""
        v = self.view()
        nc, ne, nie = v.num_cells, v.num_edges, v.num_internal_edges
        out = {
            "cell_is_owned": _as(v.cell_is_owned, nc, np.int32), "cell_local_to_owned": _as(v.cell_local_to_owned, nc, np.int32),
            "cell_global_ids": _as(v.cell_global_ids, nc, np.int64), "cell_areas": _as(v.cell_areas, nc, np.float64),
            "cell_dz_dx": _as(v.cell_dz_dx, nc, np.float64), "cell_dz_dy": _as(v.cell_dz_dy, nc, np.float64),
            "cell_xc": _as(v.cell_xc, nc, np.float64), "cell_yc": _as(v.cell_yc, nc, np.float64),
            "edge_cell_ids": _as(v.edge_cell_ids, 2 * ne, np.int32).reshape(-1, 2), "edge_lengths": _as(v.edge_lengths, ne, np.float64),
            "edge_cn": _as(v.edge_cn, ne, np.float64), "edge_sn": _as(v.edge_sn, ne, np.float64),
            "internal_edge_ids": _as(v.internal_edge_ids, nie, np.int32),
        }
        if v.cell_zc:
            out["cell_zc"] = _as(v.cell_zc, nc, np.float64)
        if v.edge_xm and v.edge_ym:
            out["edge_xm"], out["edge_ym"] = _as(v.edge_xm, ne, np.float64), _as(v.edge_ym, ne, np.float64)
        return out

    def halo_arrays(self) -> dict:
        h = self.halo()
        npeers = h.num_peers
        if npeers == 0:
            z = np.zeros(0, np.int32)
            return {"peer_ranks": z, "send_offsets": np.zeros(1, np.int32), "send_cells": z, "recv_offsets": np.zeros(1, np.int32), "recv_cells": z}
        so = _as(h.send_offsets, npeers + 1, np.int32)
        ro = _as(h.recv_offsets, npeers + 1, np.int32)
        return {"peer_ranks": _as(h.peer_ranks, npeers, np.int32), "send_offsets": so, "send_cells": _as(h.send_cells, int(so[-1]), np.int32),
                "recv_offsets": ro, "recv_cells": _as(h.recv_cells, int(ro[-1]), np.int32)}

    def close(self) -> None:
        if self._h:
            lib().rdyb_hostmesh_destroy(C.byref(self._h))
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
