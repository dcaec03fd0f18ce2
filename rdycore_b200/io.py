"""On-disk formats either side of the SWE path (include/rdyb200_io.h): Exodus II meshes and PETSc binary Vecs.

Thin ctypes layer over the C readers in csrc/io_formats.c, plus the two steps the reference performs right after reading:
building the FV mesh from the element list (DMPlex interpolation -> RDyMesh, src/rdymesh.c) and turning the side sets into
boundaries (the "Face Sets" label read in src/rdysetup.c:300-420; every remaining boundary edge goes into one auto-generated
reflecting boundary, src/rdysetup.c:422-470).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib, cases
from .mesh import HostMesh


class RdybExodusMesh(C.Structure):
    _fields_ = [
        ("num_dim", C.c_int32), ("num_nodes", C.c_int32), ("num_elems", C.c_int32), ("num_blocks", C.c_int32), ("num_side_sets", C.c_int32),
        ("x", C.POINTER(C.c_double)), ("y", C.POINTER(C.c_double)), ("z", C.POINTER(C.c_double)),
        ("elem_offsets", C.POINTER(C.c_int32)), ("elem_nodes", C.POINTER(C.c_int32)), ("elem_block", C.POINTER(C.c_int32)),
        ("block_ids", C.POINTER(C.c_int32)), ("side_set_ids", C.POINTER(C.c_int32)), ("side_set_offsets", C.POINTER(C.c_int32)),
        ("side_elems", C.POINTER(C.c_int32)), ("side_edges", C.POINTER(C.c_int32)),
    ]


def _arr(ptr, n, dtype):
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True) if n else np.zeros(0, dtype)


def read_exodus(path: str) -> dict:
    """DMPlexCreateFromFile for *.exo: dict of numpy arrays (0-based ids), see RdybExodusMesh."""
    lib = _lib.lib()
    m = RdybExodusMesh()
    _lib.check(lib.rdyb_exodus_read(str(path).encode(), C.byref(m)))
    try:
        ne, nss = m.num_elems, m.num_side_sets
        off = _arr(m.elem_offsets, ne + 1, np.int32)
        sso = _arr(m.side_set_offsets, nss + 1, np.int32)
        nsides = int(sso[-1]) if nss else 0
        return dict(
            num_dim=m.num_dim,
            xyz=np.stack([_arr(m.x, m.num_nodes, np.float64), _arr(m.y, m.num_nodes, np.float64), _arr(m.z, m.num_nodes, np.float64)], axis=1),
            elem_offsets=off, elem_nodes=_arr(m.elem_nodes, int(off[-1]), np.int32), elem_block=_arr(m.elem_block, ne, np.int32),
            block_ids=_arr(m.block_ids, m.num_blocks, np.int32), side_set_ids=_arr(m.side_set_ids, nss, np.int32), side_set_offsets=sso,
            side_elems=_arr(m.side_elems, nsides, np.int32), side_edges=_arr(m.side_edges, nsides, np.int32),
        )
    finally:
        lib.rdyb_exodus_free(C.byref(m))


def read_petsc_vec(path: str) -> np.ndarray:
    """VecLoad of a PETSc binary viewer file (32- or 64-bit index build)."""
    lib = _lib.lib()
    n, p, ib = C.c_int64(), C.POINTER(C.c_double)(), C.c_int32()
    _lib.check(lib.rdyb_petsc_vec_read(str(path).encode(), C.byref(n), C.byref(p), C.byref(ib)))
    try:
        return _arr(p, n.value, np.float64)
    finally:
        lib.rdyb_free(p)


def write_petsc_vec(path: str, values: np.ndarray, index_bytes: int = 4) -> None:
    v = np.ascontiguousarray(values, dtype=np.float64).ravel()
    _lib.check(_lib.lib().rdyb_petsc_vec_write(str(path).encode(), C.c_int64(v.size), v.ctypes.data_as(C.POINTER(C.c_double)), C.c_int32(index_bytes)))


class _CheckpointMetadata(C.Structure):
    _fields_ = [("nproc", C.c_int32), ("t", C.c_double), ("dt", C.c_double), ("step", C.c_int32)]


def checkpoint_filename_base(directory: str, prefix: str, index: int, max_index: int) -> str:
    """GenerateE3SMCheckpointFilenameBase (src/checkpoint.c:141-152); the binary format appends '.bin'."""
    buf = C.create_string_buffer(4096)
    _lib.check(_lib.lib().rdyb_checkpoint_filename_base(directory.encode(), prefix.encode(), C.c_int64(index), C.c_int64(max_index), buf, C.c_int32(4096)))
    return buf.value.decode()


def write_checkpoint(path: str, values: np.ndarray, t: float, dt: float, step: int, nproc: int = 1, num_global_cells: int | None = None,
                     natural_ids: np.ndarray | None = None, write_header: bool = True, index_bytes: int = 4) -> None:
    """WriteCheckpointFile, binary format (src/checkpoint.c:153-212): values[(local cell), component] go to the rows natural_ids of the
    natural-order solution (all rows when natural_ids is None); the rank with write_header writes the metadata bag and the Vec header."""
    v = np.ascontiguousarray(values, dtype=np.float64)
    assert v.ndim == 2
    n, ndof = v.shape
    ids = None if natural_ids is None else np.ascontiguousarray(natural_ids, dtype=np.int64)
    md = _CheckpointMetadata(nproc, t, dt, step)
    _lib.check(_lib.lib().rdyb_checkpoint_write(
        str(path).encode(), C.byref(md), C.c_int64(n if num_global_cells is None else num_global_cells), C.c_int32(ndof), C.c_int32(index_bytes), C.c_int64(n),
        None if ids is None else ids.ctypes.data_as(C.POINTER(C.c_int64)), v.ctypes.data_as(C.POINTER(C.c_double)), C.c_int32(1 if write_header else 0)))


def read_checkpoint(path: str, ndof: int = 3) -> dict:
    """ReadCheckpointFile, binary format (src/checkpoint.c:258-320): the metadata and the natural-order solution [cells, ndof]."""
    lib = _lib.lib()
    md, n, p, ib = _CheckpointMetadata(), C.c_int64(), C.POINTER(C.c_double)(), C.c_int32()
    _lib.check(lib.rdyb_checkpoint_read(str(path).encode(), C.byref(md), C.byref(n), C.byref(p), C.byref(ib)))
    try:
        u = _arr(p, n.value, np.float64)
    finally:
        lib.rdyb_free(p)
    return dict(nproc=md.nproc, t=md.t, dt=md.dt, step=md.step, index_bytes=ib.value, solution=u.reshape(-1, ndof))


def mesh_from_exodus(ex: dict) -> HostMesh:
    return HostMesh.from_cells(ex["xyz"], ex["elem_offsets"], ex["elem_nodes"])


def boundaries_from_side_sets(mesh: HostMesh, ex: dict, bc_types: dict | None = None, default_type=cases.REFLECTING):
    """One Boundary per side set (id = the deck's grid_boundary_id, type from bc_types, default reflecting) and, like the reference,
    one extra reflecting boundary holding every boundary edge no side set claims."""
    ec = mesh.view_arrays()["edge_cell_ids"]
    ev = mesh.array("edge_vertex_ids").reshape(-1, 2)
    bedges = mesh.array("boundary_edge_ids")
    key = {(int(ec[e, 0]), *sorted((int(ev[e, 0]), int(ev[e, 1])))): int(e) for e in bedges}
    off, nodes = ex["elem_offsets"], ex["elem_nodes"]
    claimed, out = set(), []
    for s, sid in enumerate(ex["side_set_ids"]):
        edges = []
        for k in range(ex["side_set_offsets"][s], ex["side_set_offsets"][s + 1]):
            el, le = int(ex["side_elems"][k]), int(ex["side_edges"][k])
            if le < 0:
                continue  # a shell face, not an edge
            n = off[el + 1] - off[el]
            a, b = int(nodes[off[el] + le]), int(nodes[off[el] + (le + 1) % n])
            e = key.get((el, *sorted((a, b))))
            if e is None:
                continue  # a side on an interior edge: the reference intersects each face set with the domain boundary (src/rdysetup.c:398)
            edges.append(e)
            claimed.add(e)
        bt = (bc_types or {}).get(int(sid), default_type)
        out.append(cases.Boundary(int(sid), np.asarray(edges, dtype=np.int32), bt, f"side_set_{int(sid)}"))
    rest = np.asarray([e for e in bedges if int(e) not in claimed], dtype=np.int32)
    if len(rest):
        unassigned_id = 1  # src/rdysetup.c:393-394: bumped past every id met in ascending order
        for bid in sorted(b.id for b in out):
            if bid == unassigned_id:
                unassigned_id += 1
        out.append(cases.Boundary(unassigned_id, rest, cases.REFLECTING, "auto_generated_reflecting"))
    return out
