"""Copies the small mesh / initial-condition / material fixtures of the reference's SWE decks (share/meshes, share/conditions,
share/materials: binary DATA files, no code) into tests/golden/ref_fixtures/ and records what an independent parser (scipy's
NetCDF-3 reader, numpy for the PETSc binary Vecs) makes of them in ref_fixtures.npz — the expected values of
tests/test_io_formats.py — together with the scalar sections of the decks (driver/tests/swe_roe/*.yaml) that use them, into tests/decks/ref/.  The tests then never touch /root/reference (it does not exist on the GPU box).
Run in the build container:  python tests/golden/make_ref_fixtures.py"""
import os
import shutil

import numpy as np
from scipy.io import netcdf_file

SHARE = "/root/reference/share"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "ref_fixtures")
MESHES = ["DamBreak_grid5x10", "DamBreak_grid5x10_mixed_elements", "quad_tri_mesh", "mms_triangles_dx1", "levee", "four_mounds_60x24", "Houston1km_with_z", "parabolic_bowl"]
VECS = [
    "conditions/DamBreak_grid5x10_wetdownstream.ic.int32.bin", "conditions/DamBreak_grid5x10_wetdownstream.ic.int64.bin",
    "conditions/DamBreak_grid5x10_mixed_elements_wetdownstream.ic.int32.bin", "materials/manning_grid5x10.int32.bin",
    "materials/manning_grid5x10_mixed_elements.int32.bin", "conditions/levee_ic.int32.bin", "meshes/levee_elevation.int32.bin",
    "conditions/four_mounds_60x24.ic.int32.bin", "conditions/parabolic_bowl.ic.int32.bin", "materials/manning_four_mound_60x24.int32.bin", "conditions/Houston1km.ic.int32.bin",
    # forcing data: homogeneous rain and boundary series, two hourly rain rasters, unstructured rain / boundary data
    "conditions/Houston1km.rain.int32.bin", "conditions/Houston1km.bc.int32.bin", "conditions/2017-08-26:00-00.int32.bin",
    "conditions/2017-08-26:01-00.int32.bin", "conditions/simons-ss/simons_forcing_x_y.int32.bin", "conditions/simons-bc/simons_bc_x_y.int32.bin",
]
RENAMED = {"conditions/simons-ss/2011-08-27:14-00.int32.bin": "simons-ss.2011-08-27:14-00.int32.bin",
           "conditions/simons-bc/2011-08-27:14-00.int32.bin": "simons-bc.2011-08-27:14-00.int32.bin"}

DECKS = ["ex2b_ic_file", "mixed_elements_ic_file", "four_mounds_60x24", "levee.hr", "Houston1km.DirichletBC"]

os.makedirs(OUT, exist_ok=True)
os.makedirs(os.path.join(HERE, "..", "decks", "ref"), exist_ok=True)
KEEP = ("physics", "numerics", "time", "grid")  # the sections RDyCreate reads here; regions / materials / conditions / output go through setters
for d in DECKS:
    lines, keep = [], False
    for ln in open(f"/root/reference/driver/tests/swe_roe/{d}.yaml"):
        if not ln.strip() or ln.lstrip().startswith("#"):
            continue
        if not ln[0].isspace():
            keep = ln.split(":")[0].strip() in KEEP
        if keep:
            lines.append(ln.rstrip() + "\n")
    with open(os.path.join(HERE, "..", "decks", "ref", d + ".yaml"), "w") as f:
        f.write(f"# physics / numerics / time / grid sections of the reference deck driver/tests/swe_roe/{d}.yaml, values unchanged\n"
                "# (written by tests/golden/make_ref_fixtures.py).  Its regions, materials, initial and boundary conditions are applied by\n"
                "# tests/test_reference_decks.py through the setters, from the fixture files the deck names.\n")
        f.writelines(lines)
expected = {}
for m in MESHES:
    src = f"{SHARE}/meshes/{m}.exo"
    shutil.copyfile(src, os.path.join(OUT, m + ".exo"))
    f = netcdf_file(src, "r", mmap=False)
    nb = f.dimensions["num_el_blk"]
    expected[m + "/xyz"] = np.stack([f.variables["coordx"][:], f.variables["coordy"][:], f.variables["coordz"][:]], axis=1).astype(np.float64)
    conn, blk = [], []
    ids = f.variables["eb_prop1"][:]
    for b in range(1, nb + 1):
        c = f.variables[f"connect{b}"][:].astype(np.int32) - 1
        conn += [row for row in c]
        blk += [int(ids[b - 1])] * len(c)
    expected[m + "/elem_offsets"] = np.cumsum([0] + [len(r) for r in conn]).astype(np.int32)
    expected[m + "/elem_nodes"] = np.concatenate(conn).astype(np.int32)
    expected[m + "/elem_block"] = np.asarray(blk, np.int32)
    nss = f.dimensions.get("num_side_sets", 0) or 0
    expected[m + "/side_set_ids"] = f.variables["ss_prop1"][:].astype(np.int32) if nss else np.zeros(0, np.int32)
    se, sr = [], []
    for s in range(1, nss + 1):
        se.append(f.variables[f"elem_ss{s}"][:].astype(np.int32) - 1)
        sr.append(f.variables[f"side_ss{s}"][:].astype(np.int32))  # RAW Exodus side numbers (1-based)
    expected[m + "/side_set_offsets"] = np.cumsum([0] + [len(x) for x in se]).astype(np.int32)
    expected[m + "/side_elems"] = np.concatenate(se).astype(np.int32) if se else np.zeros(0, np.int32)
    expected[m + "/side_raw"] = np.concatenate(sr).astype(np.int32) if sr else np.zeros(0, np.int32)
for v in VECS + list(RENAMED):
    src = f"{SHARE}/{v}"
    name = RENAMED.get(v, os.path.basename(v))
    shutil.copyfile(src, os.path.join(OUT, name))
    raw = open(src, "rb").read()
    if raw[:4] == (1211214).to_bytes(4, "big"):
        n = int.from_bytes(raw[4:8], "big"); off = 8
    else:
        assert raw[:8] == (1211214).to_bytes(8, "big")
        n = int.from_bytes(raw[8:16], "big"); off = 16
    expected["vec/" + name] = np.frombuffer(raw, dtype=">f8", count=n, offset=off).astype(np.float64)
np.savez_compressed(os.path.join(HERE, "ref_fixtures.npz"), **expected)
print({k: v.shape for k, v in expected.items()})
