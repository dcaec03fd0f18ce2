"""Named workloads (BASELINE.json configs A-E) as host-side setups: mesh, boundaries, initial state, Manning n, sources.

This plays the role of the reference's YAML decks (driver/tests/swe_roe/*.yaml) for the hot path only:
everything here ends in the operator's setters, exactly like InitBoundaries / InitFlowSolution /
InitMaterialProperties (src/rdysetup.c:283-503, 541-590, 781-879) end in operator data.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from ._lib import RdybBoundaryView, RdybConfig
from .mesh import HostMesh

DIRICHLET, NEUMANN, REFLECTING, CRITICAL_OUTFLOW = 0, 1, 2, 3
SEMI_IMPLICIT, IMPLICIT_XQ2018 = 0, 1
WB_NONE, WB_BS2002, WB_HR = 0, 1, 2
EULER, RK4 = 0, 1
GRAVITY = 9.806


def make_config(source_method=SEMI_IMPLICIT, well_balancing=WB_NONE, temporal=EULER, tiny_h=1e-7, h_anuga=0.0, xq2018_threshold=1e-10):
    """Defaults of src/yaml_input.c:854-865."""
    return RdybConfig(tiny_h, h_anuga, xq2018_threshold, source_method, well_balancing, temporal, 0)


@dataclass
class Boundary:
    id: int
    edge_ids: np.ndarray  # local edge ids, int32
    bc_type: int
    name: str = ""


@dataclass
class Case:
    name: str
    mesh: HostMesh
    config: RdybConfig
    boundaries: list
    u0: np.ndarray  # [num_owned, 3] AoS h, hu, hv
    mannings: np.ndarray  # [num_owned]
    sources: np.ndarray | None  # [num_owned, 3] or None
    dt: float
    dirichlet: dict = field(default_factory=dict)  # boundary index -> [ne, 3]
    info: dict = field(default_factory=dict)

    def boundary_views(self):
        """ctypes array of RdybBoundaryView (keeps the edge id arrays alive on self)."""
        nb = len(self.boundaries)
        arr = (RdybBoundaryView * max(nb, 1))()
        self._bkeep = []
        for i, b in enumerate(self.boundaries):
            ids = np.ascontiguousarray(b.edge_ids, dtype=np.int32)
            self._bkeep.append(ids)
            arr[i] = RdybBoundaryView(b.id, i, len(ids), ids.ctypes.data_as(C.POINTER(C.c_int32)), b.bc_type)
        self._barr = arr
        return arr, nb


def boundaries_from_predicates(mesh: HostMesh, specs, default_type=REFLECTING):
    """specs: list of (id, name, bc_type, predicate(xc, yc) -> bool mask) over boundary-edge centroids, first match wins.
    Boundary edges matched by no spec go to an auto-generated reflecting boundary, as InitBoundaries does
    (src/rdysetup.c:425-441, 768-773)."""
    be = mesh.array("boundary_edge_ids")
    # only edges whose left cell is owned belong to this rank's boundaries
    xc, yc = mesh.array("edge_xc")[be], mesh.array("edge_yc")[be]
    taken = np.zeros(len(be), dtype=bool)
    out = []
    for bid, name, bc_type, pred in specs:
        m = pred(xc, yc) & ~taken
        taken |= m
        out.append(Boundary(bid, be[m].astype(np.int32), bc_type, name))
    rest = be[~taken].astype(np.int32)
    if len(rest):
        out.append(Boundary(max([b.id for b in out], default=0) + 1, rest, default_type, "auto_reflecting"))
    return out


def dam_break_quads(nx=100, ny=100, lx=10.0, ly=5.0, source_method=IMPLICIT_XQ2018, dt=1e-3) -> Case:
    """Config A: ex2b-style dam break (driver/tests/swe_roe/ex2b.yaml) on nx x ny quads.
    h = 10 for x < lx/2, 5 otherwise; n = 0.015; top wall reflecting, bottom wall critical outflow, side walls auto-reflecting."""
    mesh = HostMesh.quads(nx, ny, lx, ly)
    eps = 1e-9 * ly
    bnd = boundaries_from_predicates(mesh, [
        (2, "top_wall", REFLECTING, lambda x, y: y > ly - eps),
        (3, "bottom_wall", CRITICAL_OUTFLOW, lambda x, y: y < eps),
    ])
    xc = mesh.array("cell_xc")
    n = mesh.num_owned_cells
    u0 = np.zeros((n, 3))
    u0[:, 0] = np.where(xc < 0.5 * lx, 10.0, 5.0)
    return Case("dam_break_quads_%dx%d" % (nx, ny), mesh, make_config(source_method=source_method), bnd, u0, np.full(n, 0.015), None, dt)
