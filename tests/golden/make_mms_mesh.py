"""Extracts the reference's MMS base mesh (share/meshes/mms_triangles_dx1.exo: NetCDF-3 Exodus II, 61 nodes, 100 TRI3,
one side set of 20 boundary sides) into tests/golden/mms_triangles_dx1.npz, so that the tests do not need the reference tree
(it does not exist on the GPU box).  Run in the build container:  python tests/golden/make_mms_mesh.py"""
import os

import numpy as np
from scipy.io import netcdf_file

SRC = "/root/reference/share/meshes/mms_triangles_dx1.exo"
HERE = os.path.dirname(os.path.abspath(__file__))

f = netcdf_file(SRC, "r", mmap=False)
xyz = np.stack([f.variables["coordx"][:], f.variables["coordy"][:], f.variables["coordz"][:]], axis=1).astype(np.float64)
conn = (f.variables["connect1"][:].astype(np.int32) - 1)  # Exodus is 1-based
elem_ss = f.variables["elem_ss1"][:].astype(np.int32) - 1
side_ss = f.variables["side_ss1"][:].astype(np.int32) - 1
np.savez_compressed(os.path.join(HERE, "mms_triangles_dx1.npz"), xyz=xyz, cells=conn, side_set_elems=elem_ss, side_set_sides=side_ss)
print(xyz.shape, conn.shape, xyz.min(0), xyz.max(0))
