"""ctypes binding of oracle/_ref/libbs2002_oracle_ref.so (oracle/bs2002_ref_driver.c): the reference's libCEED Q-functions for the SWE
flux, Bradford-Sanders well balancing and sources, driven on a serial triangle mesh.  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_ref", "libbs2002_oracle_ref.so")
_i32p, _f64p = C.POINTER(C.c_int32), C.POINTER(C.c_double)


class Bs2002Mesh(C.Structure):
    _fields_ = [("num_cells", C.c_int32), ("num_vertices", C.c_int32), ("num_edges", C.c_int32), ("num_internal_edges", C.c_int32),
                ("num_boundary_edges", C.c_int32),
                ("cell_vertex_offsets", _i32p), ("cell_vertex_ids", _i32p), ("vertex_z", _f64p), ("edge_cell_ids", _i32p), ("edge_vertex_ids", _i32p),
                ("edge_lengths", _f64p), ("edge_cn", _f64p), ("edge_sn", _f64p), ("internal_edge_ids", _i32p),
                ("cell_areas", _f64p), ("cell_dz_dx", _f64p), ("cell_dz_dy", _f64p),
                ("boundary_edge_ids", _i32p), ("boundary_edge_type", _i32p), ("boundary_values", _f64p)]


class Bs2002Params(C.Structure):
    _fields_ = [("tiny_h", C.c_double), ("h_anuga_regular", C.c_double), ("xq2018_threshold", C.c_double), ("gravity", C.c_double),
                ("source_method", C.c_int32), ("well_balancing_bs2002", C.c_int32)]


def available() -> bool:
    return os.path.exists(LIB)


def load() -> C.CDLL:
    if not os.path.exists(LIB):
        subprocess.run(["make", "-C", HERE, "bs2002"], check=True, capture_output=True)
    return C.CDLL(LIB, mode=C.RTLD_LOCAL)


class Bs2002Oracle:
    """The libCEED-semantics operator on a serial Case (rdycore_b200.cases.Case on a triangle mesh)."""

    def __init__(self, case, bs2002=True):
        self.lib = load()
        m, va = case.mesh, case.mesh.view_arrays()
        assert m.num_cells == m.num_owned_cells, "serial meshes only"
        keep = {
            "cvo": np.ascontiguousarray(m.array("cell_offsets"), np.int32), "cvi": np.ascontiguousarray(m.array("cell_vertex_ids"), np.int32),
            "vz": np.ascontiguousarray(m.vertices()[:, 2], np.float64), "eci": np.ascontiguousarray(va["edge_cell_ids"].ravel(), np.int32),
            "evi": np.ascontiguousarray(m.array("edge_vertex_ids"), np.int32), "el": np.ascontiguousarray(va["edge_lengths"]),
            "ecn": np.ascontiguousarray(va["edge_cn"]), "esn": np.ascontiguousarray(va["edge_sn"]),
            "iei": np.ascontiguousarray(va["internal_edge_ids"], np.int32), "ca": np.ascontiguousarray(va["cell_areas"]),
            "dzx": np.ascontiguousarray(va["cell_dz_dx"]), "dzy": np.ascontiguousarray(va["cell_dz_dy"]),
        }
        be, bt, bv = [], [], []
        for i, b in enumerate(case.boundaries):
            be += list(b.edge_ids)
            bt += [b.bc_type] * len(b.edge_ids)
            vals = case.dirichlet.get(i)
            bv += list(np.asarray(vals, float).reshape(-1, 3)) if vals is not None else [np.zeros(3)] * len(b.edge_ids)
        keep["be"] = np.ascontiguousarray(be, np.int32)
        keep["bt"] = np.ascontiguousarray(bt, np.int32)
        keep["bv"] = np.ascontiguousarray(np.array(bv, float).reshape(-1, 3)) if len(bv) else np.zeros((1, 3))
        self._keep = keep
        p = lambda k, t: keep[k].ctypes.data_as(t)  # noqa: E731
        self.mesh = Bs2002Mesh(m.num_cells, len(keep["vz"]), m.num_edges, m.num_internal_edges, len(keep["be"]),
                               p("cvo", _i32p), p("cvi", _i32p), p("vz", _f64p), p("eci", _i32p), p("evi", _i32p), p("el", _f64p), p("ecn", _f64p),
                               p("esn", _f64p), p("iei", _i32p), p("ca", _f64p), p("dzx", _f64p), p("dzy", _f64p), p("be", _i32p), p("bt", _i32p),
                               p("bv", _f64p))
        cfg = case.config
        self.params = Bs2002Params(cfg.tiny_h, cfg.h_anuga_regular, cfg.xq2018_threshold, 9.806, cfg.source_method, int(bs2002))
        n = m.num_cells
        self.u = np.ascontiguousarray(case.u0, np.float64).copy()
        self.mannings = np.ascontiguousarray(case.mannings, np.float64)
        self.sources = np.ascontiguousarray(case.sources if case.sources is not None else np.zeros((n, 3)), np.float64)
        self.eta = np.zeros(len(keep["vz"]))
        self.F = np.zeros((n, 3))
        self._courant = C.c_double()

    def _p(self, a):
        return a.ctypes.data_as(_f64p)

    def rhs(self, dt: float) -> np.ndarray:
        ierr = self.lib.bs2002_oracle_rhs(C.byref(self.mesh), C.byref(self.params), C.c_double(dt), self._p(self.u), self._p(self.mannings),
                                          self._p(self.sources), self._p(self.F), self._p(self.eta), C.byref(self._courant))
        assert ierr == 0
        return self.F

    def step_euler(self, dt: float, nsteps: int) -> None:
        ierr = self.lib.bs2002_oracle_step_euler(C.byref(self.mesh), C.byref(self.params), C.c_double(dt), C.c_int32(nsteps), self._p(self.u),
                                                 self._p(self.mannings), self._p(self.sources), self._p(self.eta), C.byref(self._courant))
        assert ierr == 0

    def courant(self) -> float:
        return self._courant.value
