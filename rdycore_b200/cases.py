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
FIRST_ORDER, MUSCL_MINMOD, MUSCL_NONE, MUSCL_VANLEER = 0, 1, 2, 3  # numerics.second_order + limiter
GRAVITY = 9.806


def make_config(source_method=SEMI_IMPLICIT, well_balancing=WB_NONE, temporal=EULER, tiny_h=1e-7, h_anuga=0.0, xq2018_threshold=1e-10,
                reconstruction=FIRST_ORDER):
    """Defaults of src/yaml_input.c:854-865."""
    return RdybConfig(tiny_h, h_anuga, xq2018_threshold, source_method, well_balancing, temporal, reconstruction)


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


# ---------------------------------------------------------------------------------------------------------------
# synthetic floodplain (configs C, D, E of BASELINE.json)
# ---------------------------------------------------------------------------------------------------------------
def _hash01(ids: np.ndarray, seed: int) -> np.ndarray:
    """Deterministic U(0,1) per 64-bit id (splitmix64 finaliser) — independent of the partition a cell lands in."""
    x = (ids.astype(np.uint64) + np.uint64(seed)) * np.uint64(0x9E3779B97F4A7C15)
    x ^= x >> np.uint64(30)
    x *= np.uint64(0xBF58476D1CE4E5B9)
    x ^= x >> np.uint64(27)
    x *= np.uint64(0x94D049BB133111EB)
    x ^= x >> np.uint64(31)
    return (x >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def floodplain(nx=2236, ny=2236, dx=10.0, eta0=18.0, hr=False, px=1, py=1, rank=0, source_method=SEMI_IMPLICIT, dt=None, zamp=None,
               dam_column=None) -> Case:
    """Config C (hr=False) / D (hr=True): nx x ny quads split into triangles along alternating diagonals, jittered vertices,
    bathymetry z = 20 x/L + 5 zamp sin(2 pi x/3km) cos(2 pi y/2km) + 2 U(0,1), Manning n ~ U(0.02,0.06), rain
    1e-5 (1+sin(2 pi x/5km)) (1+cos(2 pi y/5km)) m/s, lake at rest eta0, reflecting walls.  With px*py > 1 this is rank's
    block of a px x py decomposition with one layer of face-neighbour ghosts and its halo plan."""
    lx, ly = nx * dx, ny * dx
    if zamp is None:
        zamp = 1.5 if hr else 1.0
    mesh = HostMesh.floodplain(nx, ny, lx, ly, jitter=0.2, zamp=zamp, seed=12345, px=px, py=py, rank=rank)
    if hr:
        mesh.override_2d_projection()
        mesh.set_zc(None)
    n_own = mesh.num_owned_cells
    o2l = mesh.array("owned_to_local")
    gids = mesh.array("cell_global_ids")[o2l]
    xc, yc, zc = mesh.array("cell_xc")[o2l], mesh.array("cell_yc")[o2l], mesh.array("cell_zcen")[o2l]
    if hr:
        zc = mesh.view_arrays()["cell_zc"][o2l]
    mann = 0.02 + 0.04 * _hash01(gids, 34567)
    src = np.zeros((n_own, 3))
    src[:, 0] = 1e-5 * (1.0 + np.sin(2 * np.pi * xc / 5000.0)) * (1.0 + np.cos(2 * np.pi * yc / 5000.0))
    u0 = np.zeros((n_own, 3))
    u0[:, 0] = np.maximum(0.0, eta0 - zc)
    if dam_column is not None:
        # a water column in the low corner drives wet/dry fronts (config D)
        r, height = dam_column
        u0[:, 0] = np.where((xc < r) & (yc < r), np.maximum(u0[:, 0], height - zc), u0[:, 0])
        u0[:, 0] = np.maximum(u0[:, 0], 0.0)
    u0_ghost = None
    if px * py > 1:
        # the same initial state on the ghost cells (a caller that freezes ghosts, e.g. a bounded CPU sample of the mesh, starts from it)
        n_loc = mesh.num_cells
        gl = np.setdiff1d(np.arange(n_loc, dtype=np.int64), o2l.astype(np.int64), assume_unique=True)
        gx, gy = mesh.array("cell_xc")[gl], mesh.array("cell_yc")[gl]
        gz = mesh.view_arrays()["cell_zc"][gl] if hr else mesh.array("cell_zcen")[gl]
        u0_ghost = np.zeros((len(gl), 3))
        u0_ghost[:, 0] = np.maximum(0.0, eta0 - gz)
        if dam_column is not None:
            u0_ghost[:, 0] = np.maximum(np.where((gx < dam_column[0]) & (gy < dam_column[0]), np.maximum(u0_ghost[:, 0], dam_column[1] - gz), u0_ghost[:, 0]), 0.0)
    bnd = boundaries_from_predicates(mesh, [])
    # keep only boundary edges whose left cell is owned (the reference builds boundaries per rank the same way)
    if px * py > 1:
        va = mesh.view_arrays()
        for b in bnd:
            left = va["edge_cell_ids"][b.edge_ids, 0]
            b.edge_ids = b.edge_ids[va["cell_is_owned"][left] == 1]
    if dt is None:
        hmax = max(float(u0[:, 0].max()), 1.0)
        dt = 0.25 * 0.2 * dx / np.sqrt(GRAVITY * hmax)  # Courant ~ amax*L/A*dt with L/A ~ 0.3/dx*10
    cfg = make_config(source_method=source_method, well_balancing=WB_HR if hr else WB_NONE)
    return Case("floodplain_%s_%dx%dx2" % ("hr" if hr else "roe", nx, ny), mesh, cfg, bnd, u0, mann, src, float(dt),
                info={"px": px, "py": py, "rank": rank, "eta0": eta0, "dx": dx, "u0_ghost": u0_ghost})
