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


# ---- checkpoints (src/checkpoint.c, binary format) ----
def test_checkpoint_file_names_follow_the_e3sm_pattern():
    assert io.checkpoint_filename_base("checkpoints", "ex2b", 7, 1000) == "checkpoints/ex2b.rdycore.r.0007"
    assert io.checkpoint_filename_base("checkpoints/member1", "run", 12, 99) == "checkpoints/member1/run.rdycore.r.12"
    assert io.checkpoint_filename_base("d", "p", 0, 5) == "d/p.rdycore.r.0"


@pytest.mark.parametrize("index_bytes", [4, 8])
def test_checkpoint_round_trip_and_record_layout(tmp_path, index_bytes):
    rng = np.random.default_rng(3)
    u = rng.standard_normal((37, 3))
    path = str(tmp_path / "run.rdycore.r.010.bin")
    io.write_checkpoint(path, u, t=1.25, dt=0.005, step=250, nproc=4, index_bytes=index_bytes)
    ck = io.read_checkpoint(path)
    assert (ck["nproc"], ck["t"], ck["dt"], ck["step"], ck["index_bytes"]) == (4, 1.25, 0.005, 250, index_bytes)
    assert np.array_equal(ck["solution"], u)
    raw = open(path, "rb").read()
    ib = index_bytes
    geti = lambda at: int.from_bytes(raw[at:at + ib], "big", signed=True)
    # PetscBagView: classid, deprecated size, item count, name[64], help[128]
    assert geti(0) == 1211219 and geti(ib) == 0 and geti(2 * ib) == 4
    assert raw[3 * ib:3 * ib + 64].rstrip(b"\0") == b"metadata" and raw[3 * ib + 64:3 * ib + 192].rstrip(b"\0") == b"Checkpoint metadata"
    item = 3 * ib + 192
    at = item
    for name, dtype, width in (("nproc", 16, ib), ("t", 1, 8), ("dt", 1, 8), ("step", 16, ib)):
        assert geti(at + ib) == dtype and raw[at + 2 * ib:at + 2 * ib + 64].rstrip(b"\0") == name.encode() and geti(at + 2 * ib + 192) == 1
        at += item + width
    # then the Vec record exactly as rdyb_petsc_vec_write lays it out: the tail of the file is a loadable PETSc Vec
    assert geti(at) == 1211214 and geti(at + ib) == u.size
    vec = str(tmp_path / "tail.bin")
    open(vec, "wb").write(raw[at:])
    assert np.array_equal(io.read_petsc_vec(vec), u.ravel())
    assert open(path + ".info").read() == "-vecload_block_size 3\n"


def test_checkpoint_written_by_parts_equals_the_serial_file(tmp_path):
    """One process per GPU writes its owned rows at their natural offsets; the file is the serial one, in any order of the calls."""
    rng = np.random.default_rng(4)
    n = 1000
    u = rng.standard_normal((n, 3))
    serial, parts = str(tmp_path / "a.bin"), str(tmp_path / "b.bin")
    io.write_checkpoint(serial, u, t=3.0, dt=0.1, step=30, nproc=3)
    owner = rng.integers(0, 3, n)
    open(parts, "wb").write(b"\xff" * 50000)  # a stale, longer file under the same name
    for rank in (2, 0, 1):
        ids = np.flatnonzero(owner == rank)
        ids = ids[rng.permutation(ids.size)] if rank == 1 else ids  # one rank in a scrambled (device) order
        io.write_checkpoint(parts, u[ids], t=3.0, dt=0.1, step=30, nproc=3, num_global_cells=n, natural_ids=ids, write_header=rank == 0)
    assert open(parts, "rb").read() == open(serial, "rb").read()


def test_checkpoint_errors(tmp_path):
    with pytest.raises(RuntimeError, match="Invalid checkpoint file"):
        io.read_checkpoint(str(tmp_path / "x.h5"))
    p = str(tmp_path / "v.bin")
    io.write_petsc_vec(p, np.arange(6.0))
    with pytest.raises(RuntimeError, match="metadata bag"):
        io.read_checkpoint(p)
    with pytest.raises(RuntimeError, match="out of range"):
        io.write_checkpoint(str(tmp_path / "w.bin"), np.zeros((2, 3)), 0.0, 1.0, 0, num_global_cells=2, natural_ids=np.array([0, 5]))


# ---- corrupt input: the readers answer with an error, never a crash (mutation fuzz, scratch/fuzz/fuzz_readers.py runs it at length) ----
@pytest.mark.parametrize("fixture,kind,seed", [
    ("mms_triangles_dx1.exo", "exo", 0),   # crashed before the header sanity checks: element blocks larger than num_elem
    ("quad_tri_mesh.exo", "exo", 2),       # crashed: dimension id >= 2^31 in a variable's dimension list
    ("levee.exo", "exo", 5),
    ("DamBreak_grid5x10_wetdownstream.ic.int64.bin", "vec", 1),
    ("checkpoint", "ck", 3),
])
def test_mutated_files_never_crash_the_readers(tmp_path, fixture, kind, seed):
    import subprocess
    import sys

    src = os.path.join(FIX, fixture)
    if kind == "ck":
        src = str(tmp_path / "src.rdycore.r.bin")
        io.write_checkpoint(src, np.arange(90.0).reshape(30, 3), t=1.0, dt=0.1, step=10, nproc=2, index_bytes=8)
    tmp = str(tmp_path / ("mutant.bin" if kind != "exo" else "mutant.exo"))
    r = subprocess.run([sys.executable, os.path.join(os.path.dirname(__file__), "reader_fuzz.py"), src, kind, str(seed), "50", tmp],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, f"reader crashed (rc {r.returncode}) after trial {r.stdout.strip().splitlines()[-1:]}: {r.stderr[-300:]}"
    ok, err = map(int, r.stdout.strip().splitlines()[-1].split()[1:])
    assert ok + err == 50 and err > 0
