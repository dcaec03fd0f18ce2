"""ctypes binding of the CPU oracles (oracle/swe_oracle.h).  TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs,
never by the product package rdycore_b200.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libswe_oracle_ref.so")
PORT_LIB = os.path.join(HERE, "libswe_oracle_port.so")
# NOT an oracle: the reference's driver code with the B200 adapter's factories (oracle/Makefile target `adapter`); the object under
# test of tests/test_gpu_rdycore_adapter.py, which drives it through this same binding
ADAPTER_LIB = os.path.join(HERE, "_ref", "libswe_adapter_ref.so")


class Courant(C.Structure):
    _fields_ = [("max_courant_num", C.c_double), ("global_edge_id", C.c_int64), ("global_cell_id", C.c_int64)]


def build(kind: str = "all") -> None:
    targets = {"all": ["port", "ref"], "port": ["port"], "reference": ["ref"], "adapter": ["adapter"]}[kind]
    subprocess.run(["make", "-C", HERE, *targets], check=True, capture_output=True)


def available(kind: str) -> bool:
    return os.path.exists({"reference": REF_LIB, "adapter": ADAPTER_LIB}.get(kind, PORT_LIB))


_libs: dict = {}


def load(kind: str) -> C.CDLL:
    """kind = 'reference' (the reference's own code, oracle/_ref) or 'port' (restatement)."""
    if kind not in _libs:
        path = {"reference": REF_LIB, "adapter": ADAPTER_LIB}.get(kind, PORT_LIB)
        if not os.path.exists(path):
            build(kind if kind in ("reference", "adapter") else "port")
        lib = C.CDLL(path, mode=C.RTLD_LOCAL)
        lib.swe_oracle_kind.restype = C.c_char_p
        lib.swe_oracle_array.restype = C.POINTER(C.c_double)
        lib.swe_oracle_array.argtypes = [C.c_void_p, C.c_char_p, C.c_int32, C.POINTER(C.c_int64)]
        lib.swe_oracle_destroy.argtypes = [C.c_void_p]
        lib.swe_oracle_destroy.restype = None
        for fn in ("swe_oracle_rhs", "swe_oracle_euler_update"):
            getattr(lib, fn).argtypes = [C.c_void_p, C.c_double]
        for fn in ("swe_oracle_step_euler", "swe_oracle_step_rk4"):
            getattr(lib, fn).argtypes = [C.c_void_p, C.c_double, C.c_int32]
        lib.swe_oracle_courant.argtypes = [C.c_void_p, C.POINTER(Courant)]
        assert lib.swe_oracle_kind().decode() == kind
        _libs[kind] = lib
    return _libs[kind]


def best_kind() -> str:
    return "reference" if available("reference") or os.path.isdir("/root/reference/src/swe") else "port"


class Oracle:
    """One rank's Operator + solution Vecs, CPU."""

    def __init__(self, kind: str, config, mesh_view, boundaries):
        """config: RdybConfig; mesh_view: RdybMeshView (kept alive by the caller); boundaries: ctypes array of RdybBoundaryView."""
        self.kind = kind
        self.lib = load(kind)
        self._keep = (config, mesh_view, boundaries)
        self._h = C.c_void_p()
        nb = len(boundaries) if boundaries is not None else 0
        ierr = self.lib.swe_oracle_create(C.byref(config), C.byref(mesh_view), C.c_int32(nb), boundaries, C.byref(self._h))
        if ierr:
            raise RuntimeError(f"swe_oracle_create failed with code {ierr}")
        self.num_owned = mesh_view.num_owned_cells
        self.num_local = mesh_view.num_cells

    def array(self, name: str, boundary: int = -1) -> np.ndarray:
        n = C.c_int64()
        p = self.lib.swe_oracle_array(self._h, name.encode(), boundary, C.byref(n))
        if not p:
            raise KeyError(name)
        if n.value == 0:
            return np.zeros(0)
        return np.ctypeslib.as_array(p, shape=(n.value,))

    def rhs(self, dt: float) -> np.ndarray:
        ierr = self.lib.swe_oracle_rhs(self._h, dt)
        if ierr:
            raise RuntimeError(f"oracle rhs failed: {ierr}")
        return self.array("rhs")

    def euler_update(self, dt: float) -> None:
        self.lib.swe_oracle_euler_update(self._h, dt)

    def step_euler(self, dt: float, nsteps: int) -> None:
        ierr = self.lib.swe_oracle_step_euler(self._h, dt, nsteps)
        if ierr:
            raise RuntimeError(f"oracle step failed: {ierr}")

    def step_rk4(self, dt: float, nsteps: int) -> None:
        ierr = self.lib.swe_oracle_step_rk4(self._h, dt, nsteps)
        if ierr:
            raise RuntimeError(f"oracle rk4 step failed: {ierr}")

    def courant(self) -> float:
        c = Courant()
        self.lib.swe_oracle_courant(self._h, C.byref(c))
        return c.max_courant_num

    def close(self):
        if self._h:
            self.lib.swe_oracle_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
