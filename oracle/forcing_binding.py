"""ctypes binding of the forcing-data oracles (oracle/forcing_oracle.h).  TEST INFRASTRUCTURE ONLY (tests/ may import it, the
product package never does)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import binding

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libforcing_oracle_ref.so")
PORT_LIB = os.path.join(HERE, "libforcing_oracle_port.so")
_D, _I = C.POINTER(C.c_double), C.POINTER(C.c_int32)


def available(kind: str) -> bool:
    return os.path.exists(REF_LIB if kind == "reference" else PORT_LIB)


def best_kind() -> str:
    return "reference" if available("reference") else "port"


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_D)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_I)


class ForcingOracle:
    def __init__(self, kind: str | None = None):
        kind = kind or best_kind()
        path = REF_LIB if kind == "reference" else PORT_LIB
        if not os.path.exists(path):
            binding.build("reference" if kind == "reference" else "port")
        self.lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        self.lib.forcing_oracle_kind.restype = C.c_char_p
        assert self.lib.forcing_oracle_kind().decode() == kind
        self.kind = kind

    def current_data(self, pairs, t, interpolate):
        p, pp = _d(pairs)
        idx, val = C.c_int32(-1), C.c_double()
        assert self.lib.forcing_oracle_current_data(pp, C.c_int32(p.size // 2), C.c_double(t), C.c_int32(int(interpolate)), C.byref(idx), C.byref(val)) == 0
        return idx.value, val.value

    def map_unstructured(self, mx, my, dx, dy):
        (mx, pmx), (my, pmy), (dx, pdx), (dy, pdy) = _d(mx), _d(my), _d(dx), _d(dy)
        out = np.zeros(mx.size, np.int32)
        assert self.lib.forcing_oracle_map_unstructured(C.c_int32(mx.size), pmx, pmy, C.c_int32(dx.size), pdx, pdy, out.ctypes.data_as(_I)) == 0
        return out

    def map_raster(self, mx, my, ncols, nrows, xlc, ylc, cellsize):
        (mx, pmx), (my, pmy) = _d(mx), _d(my)
        out = np.zeros(mx.size, np.int32)
        assert self.lib.forcing_oracle_map_raster(C.c_int32(mx.size), pmx, pmy, C.c_int32(ncols), C.c_int32(nrows), C.c_double(xlc), C.c_double(ylc),
                                                  C.c_double(cellsize), out.ctypes.data_as(_I)) == 0
        return out

    def set_raster(self, vec, header_offset, idx):
        (v, pv), (ix, pi) = _d(vec), _i(idx)
        out = np.zeros(ix.size)
        assert self.lib.forcing_oracle_set_raster(pv, C.c_int32(v.size), C.c_int32(header_offset), C.c_int32(ix.size), pi, out.ctypes.data_as(_D)) == 0
        return out

    def set_unstructured(self, vec, stride, idx):
        (v, pv), (ix, pi) = _d(vec), _i(idx)
        out = np.zeros(ix.size * stride)
        assert self.lib.forcing_oracle_set_unstructured(pv, C.c_int32(v.size), C.c_int32(stride), C.c_int32(ix.size), pi, out.ctypes.data_as(_D)) == 0
        return out

    def homogeneous(self, pairs, t, interpolate, boundary, values, state):
        """values: float64 array updated in place ([n] or [3n]); state: [cur_idx, prev_idx] list updated in place"""
        p, pp = _d(pairs)
        cur, prev = C.c_int32(state[0]), C.c_int32(state[1])
        n = values.size // (3 if boundary else 1)
        assert self.lib.forcing_oracle_homogeneous(pp, C.c_int32(p.size // 2), C.c_double(t), C.c_int32(int(interpolate)), C.c_int32(int(boundary)), C.c_int32(n),
                                                   values.ctypes.data_as(_D), C.byref(cur), C.byref(prev)) == 0
        state[0], state[1] = cur.value, prev.value
        return values
