"""ctypes mirror of include/rdyb200_forcing.h: forcing / boundary datasets applied on the device.  No arithmetic in Python."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import check, lib

MM_PER_HR_TO_M_PER_S = 1.0 / (1000.0 * 3600.0)  # src/forcing/rdyforcing_dataset.c:303
_D, _I = C.POINTER(C.c_double), C.POINTER(C.c_int32)


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_D)


def current_data(pairs, cur_time: float, interpolate: bool):
    """RDyForcingGetCurrentData: (index of the bracket, value)"""
    p, pp = _d(pairs)
    idx, val = C.c_int32(), C.c_double()
    check(lib().rdyb_forcing_current_data(pp, C.c_int32(p.size // 2), C.c_double(cur_time), C.c_int32(int(interpolate)), C.byref(idx), C.byref(val)))
    return idx.value, val.value


class DatasetMap:
    """nearest-data-point index of every owned cell / boundary edge, resident on the operator's GPU"""

    def __init__(self, handle, n):
        self._h, self.n = handle, n

    @classmethod
    def unstructured(cls, op, mesh_xc, mesh_yc, data_xc, data_yc):
        (mx, pmx), (my, pmy), (dx, pdx), (dy, pdy) = _d(mesh_xc), _d(mesh_yc), _d(data_xc), _d(data_yc)
        h = C.c_void_p()
        check(lib().rdyb_dataset_map_create_unstructured(op._h, C.c_int32(mx.size), pmx, pmy, C.c_int32(dx.size), pdx, pdy, C.byref(h)))
        return cls(h, mx.size)

    @classmethod
    def raster(cls, op, mesh_xc, mesh_yc, ncols, nrows, xlc, ylc, cellsize):
        (mx, pmx), (my, pmy) = _d(mesh_xc), _d(mesh_yc)
        h = C.c_void_p()
        check(lib().rdyb_dataset_map_create_raster(op._h, C.c_int32(mx.size), pmx, pmy, C.c_int32(ncols), C.c_int32(nrows), C.c_double(xlc), C.c_double(ylc),
                                                   C.c_double(cellsize), C.byref(h)))
        return cls(h, mx.size)

    @classmethod
    def from_indices(cls, op, idx, ndata):
        ix = np.ascontiguousarray(idx, dtype=np.int32)
        h = C.c_void_p()
        check(lib().rdyb_dataset_map_create_from_indices(op._h, C.c_int32(ix.size), ix.ctypes.data_as(_I), C.c_int32(ndata), C.byref(h)))
        return cls(h, ix.size)

    def indices(self) -> np.ndarray:
        out = np.zeros(self.n, np.int32)
        check(lib().rdyb_dataset_map_get(self._h, C.c_int32(self.n), out.ctypes.data_as(_I)))
        return out

    def close(self):
        if self._h:
            lib().rdyb_dataset_map_destroy(C.byref(self._h))
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def set_source_from_dataset(op, component, dmap: DatasetMap, data, offset, stride=1, scale=1.0):
    v, pv = _d(data)
    check(lib().rdyb_operator_set_source_from_dataset(op._h, C.c_int32(component), dmap._h, pv, C.c_int32(v.size), C.c_int32(offset), C.c_int32(stride), C.c_double(scale)))


def set_boundary_values_from_dataset(op, b, dmap: DatasetMap, data, offset=2):
    v, pv = _d(data)
    check(lib().rdyb_operator_set_boundary_values_from_dataset(op._h, C.c_int32(b), dmap._h, pv, C.c_int32(v.size), C.c_int32(offset)))


def fill_source(op, component, value):
    check(lib().rdyb_operator_fill_source(op._h, C.c_int32(component), C.c_double(value)))


def fill_boundary_values(op, b, h, hu=0.0, hv=0.0):
    check(lib().rdyb_operator_fill_boundary_values(op._h, C.c_int32(b), C.c_double(h), C.c_double(hu), C.c_double(hv)))


def get_boundary_values(op, b, num_edges) -> np.ndarray:
    out = np.zeros((num_edges, 3))
    check(lib().rdyb_operator_get_boundary_values(op._h, C.c_int32(b), C.c_int32(num_edges), out.ctypes.data_as(_D)))
    return out


def get_state_at(op, owned_cell_ids) -> np.ndarray:
    """(h, hu, hv) at observation sites: [n, 3]"""
    ids = np.ascontiguousarray(owned_cell_ids, dtype=np.int32)
    out = np.zeros((ids.size, 3))
    check(lib().rdyb_operator_get_state_at(op._h, C.c_int32(ids.size), ids.ctypes.data_as(_I), out.ctypes.data_as(_D)))
    return out


def reset_boundary_fluxes(op):
    check(lib().rdyb_operator_reset_boundary_fluxes(op._h))
