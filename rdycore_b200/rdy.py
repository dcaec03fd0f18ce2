"""ctypes mirror of include/rdycore_b200.h: RDycore's public lifecycle (include/rdycore.h) on the B200 SWE path.

Function names are the reference's (RDyCreate, RDySetup, RDyAdvance, RDyFinished, RDyGetOwnedCellHeights ...), so
a driver written against rdycore.h (driver/main.c:40-140) reads the same here.  No arithmetic happens in Python.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import check, lib
from .petsc_operator import Vec

TIME_UNSET, TIME_SECONDS, TIME_MINUTES, TIME_HOURS, TIME_DAYS, TIME_MONTHS, TIME_YEARS = range(7)


def RDyInit():
    check(lib().RDyInit(0, None, None))


def RDyFinalize():
    lib().RDyFinalize()


def _f64(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


class RDy:
    def __init__(self, config_file: str, comm: int = 0):
        """RDyCreate(comm, config_file, &rdy)"""
        self._h = C.c_void_p()
        self._keep = []
        check(lib().RDyCreate(C.c_int(comm), config_file.encode(), C.byref(self._h)))

    # ---- what replaces DMPlex + the list sections of the deck ----
    def set_mesh(self, mesh, boundary_views=None, num_boundaries=0):
        """RDyB200SetMesh: mesh = rdycore_b200.mesh.HostMesh; boundary_views = ctypes array of RdybBoundaryView"""
        self._keep += [mesh, boundary_views]
        check(lib().RDyB200SetMesh(self._h, mesh._h, C.c_int32(num_boundaries), boundary_views))

    def add_region(self, owned_cell_ids) -> int:
        ids = np.ascontiguousarray(owned_cell_ids, dtype=np.int32)
        idx = C.c_int32()
        check(lib().RDyB200AddRegion(self._h, C.c_int32(len(ids)), ids.ctypes.data_as(C.POINTER(C.c_int32)), C.byref(idx)))
        return idx.value

    def set_device(self, device: int):
        check(lib().RDyB200SetDevice(self._h, C.c_int(device)))

    def set_kernel(self, kind: int, tile_cells: int):
        check(lib().RDyB200SetKernel(self._h, C.c_int32(kind), C.c_int32(tile_cells)))

    def max_courant_number(self) -> float:
        v = C.c_double()
        check(lib().RDyB200GetMaxCourantNumber(self._h, C.byref(v)))
        return v.value

    # ---- rdycore.h ----
    def RDySetup(self):
        check(lib().RDySetup(self._h))

    def RDyAdvance(self):
        check(lib().RDyAdvance(self._h))

    def RDyFinished(self) -> bool:
        return bool(lib().RDyFinished(self._h))

    def RDyDestroy(self):
        if self._h:
            check(lib().RDyDestroy(C.byref(self._h)))
            self._h = C.c_void_p()

    def RDyGetTime(self, unit=TIME_SECONDS) -> float:
        v = C.c_double()
        check(lib().RDyGetTime(self._h, C.c_int(unit), C.byref(v)))
        return v.value

    def RDyGetTimeStep(self, unit=TIME_SECONDS) -> float:
        v = C.c_double()
        check(lib().RDyGetTimeStep(self._h, C.c_int(unit), C.byref(v)))
        return v.value

    def RDyGetTimeUnit(self) -> int:
        v = C.c_int()
        check(lib().RDyGetTimeUnit(self._h, C.byref(v)))
        return v.value

    def RDyGetStep(self) -> int:
        v = C.c_int32()
        check(lib().RDyGetStep(self._h, C.byref(v)))
        return v.value

    def RDyGetCouplingInterval(self, unit=TIME_SECONDS) -> float:
        v = C.c_double()
        check(lib().RDyGetCouplingInterval(self._h, C.c_int(unit), C.byref(v)))
        return v.value

    def RDySetCouplingInterval(self, unit, interval: float):
        check(lib().RDySetCouplingInterval(self._h, C.c_int(unit), C.c_double(interval)))

    def _int_getter(self, fn, *args) -> int:
        v = C.c_int32()
        check(getattr(lib(), fn)(self._h, *args, C.byref(v)))
        return v.value

    def RDyGetNumOwnedCells(self) -> int:
        return self._int_getter("RDyGetNumOwnedCells")

    def RDyGetNumGlobalCells(self) -> int:
        return self._int_getter("RDyGetNumGlobalCells")

    def RDyGetNumBoundaryConditions(self) -> int:
        return self._int_getter("RDyGetNumBoundaryConditions")

    def RDyGetNumBoundaryEdges(self, b: int) -> int:
        return self._int_getter("RDyGetNumBoundaryEdges", C.c_int32(b))

    def RDyGetBoundaryConditionFlowType(self, b: int) -> int:
        return self._int_getter("RDyGetBoundaryConditionFlowType", C.c_int32(b))

    def RDyGetBoundaryID(self, b: int) -> int:
        return self._int_getter("RDyGetBoundaryID", C.c_int32(b))

    def _owned_getter(self, fn, size=None) -> np.ndarray:
        n = self.RDyGetNumOwnedCells() if size is None else size
        out = np.zeros(n)
        check(getattr(lib(), fn)(self._h, C.c_int32(n), out.ctypes.data_as(C.POINTER(C.c_double))))
        return out

    def RDyGetOwnedCellHeights(self, size=None):
        return self._owned_getter("RDyGetOwnedCellHeights", size)

    def RDyGetOwnedCellXMomenta(self, size=None):
        return self._owned_getter("RDyGetOwnedCellXMomenta", size)

    def RDyGetOwnedCellYMomenta(self, size=None):
        return self._owned_getter("RDyGetOwnedCellYMomenta", size)

    def RDyGetOwnedCellXCentroids(self):
        return self._owned_getter("RDyGetOwnedCellXCentroids")

    def RDyGetOwnedCellYCentroids(self):
        return self._owned_getter("RDyGetOwnedCellYCentroids")

    def RDyGetOwnedCellAreas(self):
        return self._owned_getter("RDyGetOwnedCellAreas")

    def RDySetFlowDirichletBoundaryValues(self, b: int, num_edges: int, ndof: int, values):
        v, p = _f64(values)
        check(lib().RDySetFlowDirichletBoundaryValues(self._h, C.c_int32(b), C.c_int32(num_edges), C.c_int32(ndof), p))

    def _set_owned(self, fn, values, *args):
        v, p = _f64(values)
        check(getattr(lib(), fn)(self._h, *args, C.c_int32(v.shape[0]), p))

    def RDySetDomainWaterSource(self, values):
        self._set_owned("RDySetDomainWaterSource", values)

    def RDySetDomainXMomentumSource(self, values):
        self._set_owned("RDySetDomainXMomentumSource", values)

    def RDySetDomainYMomentumSource(self, values):
        self._set_owned("RDySetDomainYMomentumSource", values)

    def RDySetRegionalWaterSource(self, region: int, values):
        self._set_owned("RDySetRegionalWaterSource", values, C.c_int32(region))

    def RDySetDomainManningsN(self, values):
        self._set_owned("RDySetDomainManningsN", values)

    def RDySetRegionalManningsN(self, region: int, values):
        self._set_owned("RDySetRegionalManningsN", values, C.c_int32(region))

    def RDySetInitialConditions(self, u0):
        """u0: [num_owned, 3] (h, hu, hv): goes through a prognostic Vec as in the reference (driver/main.c)."""
        h = C.c_void_p()
        check(lib().RDyCreatePrognosticVec(self._h, C.byref(h)))
        vec = Vec.__new__(Vec)
        vec._h, vec.local_size, vec.block_size = h, 3 * self.RDyGetNumOwnedCells(), 3
        vec.set(u0)
        check(lib().RDySetInitialConditions(self._h, vec.handle))
        vec.destroy()

    def RDyB200WriteCheckpoint(self, filename_base: str, nproc: int = 1, rank: int = 0, num_global_cells: int | None = None) -> str:
        """WriteCheckpointFile, binary format (src/checkpoint.c:153-212); returns the file name."""
        n = self.RDyGetNumOwnedCells() if num_global_cells is None else num_global_cells
        check(lib().RDyB200WriteCheckpoint(self._h, filename_base.encode(), C.c_int32(nproc), C.c_int32(rank), C.c_int32(n)))
        return filename_base + ".bin"

    def RDyB200ReadCheckpoint(self, filename: str, reinitialize: bool = False):
        """ReadCheckpointFile (src/checkpoint.c:258-320): state, dt, and (unless reinitialize) time and step."""
        check(lib().RDyB200ReadCheckpoint(self._h, filename.encode(), C.c_int(1 if reinitialize else 0)))

    def state(self) -> np.ndarray:
        return np.stack([self.RDyGetOwnedCellHeights(), self.RDyGetOwnedCellXMomenta(), self.RDyGetOwnedCellYMomenta()], axis=1)
