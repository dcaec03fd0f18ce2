"""CPU: the C readers of include/rdyb200_io.h (Exodus II over NetCDF-3, PETSc binary Vec) against an independent parse of the
reference's own fixture files.  Files and expected values: tests/golden/ref_fixtures/ + ref_fixtures.npz, made from
/root/reference/share by tests/golden/make_ref_fixtures.py (scipy's NetCDF reader, numpy big-endian view) — bit-exact bar:
these are integers and verbatim doubles."""
import os

import numpy as np
import pytest

from rdycore_b200 import _lib, io

HERE = os.path.dirname(os.path.abspath(__file__))
FIX = os.path.join(HERE, "golden", "ref_fixtures")
EXPECT = np.load(os.path.join(HERE, "golden", "ref_fixtures.npz"))
MESHES = sorted({k.split("/")[0] for k in EXPECT.files if not k.startswith("vec/")})
VECS = sorted(k[4:] for k in EXPECT.files if k.startswith("vec/"))


@pytest.mark.parametrize("name", MESHES)
def test_exodus_reader_matches_independent_parse(name):
    ex = io.read_exodus(os.path.join(FIX, name + ".exo"))
    for key in ("xyz", "elem_offsets", "elem_nodes", "elem_block", "side_set_ids", "side_set_offsets", "side_elems"):
        want = EXPECT[f"{name}/{key}"]
        assert ex[key].shape == want.shape and np.array_equal(ex[key], want), key
    # side numbers: 2-D elements 1..n, shells (SHELL4, TRI3 in a 3-D file) 3..n+2 -> local edge 0..n-1
    raw = EXPECT[f"{name}/side_raw"]
    n_per = np.diff(ex["elem_offsets"])[ex["side_elems"]] if len(raw) else np.zeros(0, int)
    assert ((ex["side_edges"] >= 0) & (ex["side_edges"] < n_per)).all()
    assert np.array_equal(ex["side_edges"], raw - 3)  # every fixture with side sets is a 3-D file of shell-type elements


@pytest.mark.parametrize("name", VECS)
def test_petsc_vec_reader_matches_independent_parse(name):
    v = io.read_petsc_vec(os.path.join(FIX, name))
    assert np.array_equal(v, EXPECT["vec/" + name])


def test_petsc_vec_32_and_64_bit_headers_hold_the_same_values(tmp_path):
    a = io.read_petsc_vec(os.path.join(FIX, "DamBreak_grid5x10_wetdownstream.ic.int32.bin"))
    b = io.read_petsc_vec(os.path.join(FIX, "DamBreak_grid5x10_wetdownstream.ic.int64.bin"))
    assert np.array_equal(a, b) and a.size == 3 * 44
    for ib in (4, 8):  # write -> read round trip, and byte-identical to the reference's own file for the 32-bit build
        p = tmp_path / f"roundtrip{ib}.bin"
        io.write_petsc_vec(p, a, ib)
        assert np.array_equal(io.read_petsc_vec(p), a)
    assert open(tmp_path / "roundtrip4.bin", "rb").read() == open(os.path.join(FIX, "DamBreak_grid5x10_wetdownstream.ic.int32.bin"), "rb").read()


def test_readers_fail_loudly(tmp_path):
    with pytest.raises(_lib.RdybError, match="Cannot open"):
        io.read_exodus(tmp_path / "missing.exo")
    bad = tmp_path / "bad.exo"
    bad.write_bytes(b"\x89HDF\r\n\x1a\n" + b"\0" * 64)
    with pytest.raises(_lib.RdybError, match="not a NetCDF-3"):
        io.read_exodus(bad)
    with pytest.raises(_lib.RdybError, match="not a PETSc binary Vec"):
        io.read_petsc_vec(bad)
    trunc = tmp_path / "trunc.exo"
    trunc.write_bytes(open(os.path.join(FIX, "levee.exo"), "rb").read()[:900])
    with pytest.raises(_lib.RdybError):
        io.read_exodus(trunc)


def test_mesh_and_boundaries_from_the_reference_meshes():
    """Element list -> FV mesh (areas, edges) and side sets -> boundaries, on the meshes that carry side sets."""
    ex = io.read_exodus(os.path.join(FIX, "quad_tri_mesh.exo"))
    mesh = io.mesh_from_exodus(ex)
    assert mesh.num_owned_cells == 12 and mesh.num_vertices == 14
    bnd = io.boundaries_from_side_sets(mesh, ex)
    # side set 5 lists 5 sides, some of them on interior edges: only its domain-boundary edges make the boundary
    assert [b.id for b in bnd[:5]] == [1, 2, 3, 4, 5] and [len(b.edge_ids) for b in bnd[:4]] == [2, 2, 3, 3] and len(bnd[4].edge_ids) <= 5
    all_edges = np.concatenate([b.edge_ids for b in bnd])
    assert len(set(all_edges.tolist())) == len(all_edges) == mesh.num_boundary_edges  # every boundary edge in exactly one boundary
    ex = io.read_exodus(os.path.join(FIX, "mms_triangles_dx1.exo"))
    mesh = io.mesh_from_exodus(ex)
    bnd = io.boundaries_from_side_sets(mesh, ex)
    assert len(bnd) == 1 and len(bnd[0].edge_ids) == 20 == mesh.num_boundary_edges
    assert np.isclose(mesh.view_arrays()["cell_areas"][:100].sum(), 25.0)
