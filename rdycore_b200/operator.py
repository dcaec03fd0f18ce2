"""Host-side handle of the B200 SWE operator (include/rdyb200.h) — the Python mirror of RDycore's Operator.

Mirrors the reference's Operator interface for the SWE/PETSc case: CreateOperator (src/operator.c:348-417),
ApplyOperator (:680-690), SetOperatorBoundaryValues (:1045-1061), Get/RestoreOperator*ExternalSource /
*MaterialProperties (:1203-1290, 1394-1464), Reset/GetOperatorDiagnostics (:757-768, 867-893).  All compute goes
through the C-ABI into CUDA; there is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import RdybCourantDiagnostics, RdybVertexView, check, lib

GATHER, TILED, TILED_ASYNC, TILED_FIXED = 0, 1, 2, 3


def _f64(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.c_void_p)


class Operator:
    def __init__(self, config, mesh, boundaries=None, num_boundaries=0, halo=None, device=0):
        """config: RdybConfig; mesh: HostMesh (kept alive); boundaries: ctypes array of RdybBoundaryView."""
        self._keep = (config, mesh, boundaries)
        self._view = mesh.view()
        self._halo = halo
        self._h = C.c_void_p()
        halo_ref = C.byref(halo) if halo is not None else None
        check(lib().rdyb_operator_create(C.byref(config), C.byref(self._view), C.c_int32(num_boundaries), boundaries, halo_ref,
                                         C.c_int(device), None, C.byref(self._h)))
        n = C.c_int32()
        check(lib().rdyb_operator_num_owned_cells(self._h, C.byref(n)))
        self.num_owned = n.value
        check(lib().rdyb_operator_num_local_cells(self._h, C.byref(n)))
        self.num_local = n.value

    @classmethod
    def from_case(cls, case, device=0, halo=None):
        arr, nb = case.boundary_views()
        op = cls(case.config, case.mesh, arr, nb, halo=halo, device=device)
        op.set_mannings(case.mannings)
        if case.sources is not None:
            for c in range(3):
                if c == 0 or np.any(case.sources[:, c] != 0.0):
                    op.set_external_source(c, case.sources[:, c])
        for b, vals in case.dirichlet.items():
            op.set_boundary_values(b, vals)
        op.set_state(case.u0)
        return op

    # ---- options ----
    def set_options(self, courant=True, flux_divergence=False, primitive_variables=False, boundary_fluxes=False):
        check(lib().rdyb_operator_set_options(self._h, int(courant), int(flux_divergence), int(primitive_variables), int(boundary_fluxes)))

    def set_kernel(self, kind: int, tile_cells: int = 512):
        check(lib().rdyb_operator_set_kernel(self._h, C.c_int32(kind), C.c_int32(tile_cells)))

    def set_max_ctas(self, n: int):
        """Cap the grid of the persistent kernels: every CTA then walks many tiles (the bench regime on a small mesh)."""
        check(lib().rdyb_operator_set_max_ctas(self._h, C.c_int32(n)))

    def wait_cycles(self):
        """(sum, longest, count) of the in-kernel neighbour waits since the last call, in clock64 ticks."""
        s, m, n = C.c_uint64(), C.c_uint64(), C.c_uint64()
        check(lib().rdyb_operator_get_wait_cycles(self._h, C.byref(s), C.byref(m), C.byref(n)))
        return s.value, m.value, n.value

    # ---- data ----
    def set_boundary_values(self, boundary_index: int, values, comp_offset: int = 0):
        """values: [num_edges, num_comp] (strided values[num_comp*e + c] as SetOperatorBoundaryValues takes them)."""
        v = np.ascontiguousarray(values, dtype=np.float64)
        if v.ndim == 1:
            v = v.reshape(-1, 1)
        check(lib().rdyb_operator_set_boundary_values(self._h, C.c_int32(boundary_index), C.c_int32(comp_offset), C.c_int32(v.shape[1]),
                                                      C.c_int32(v.shape[0]), v.ctypes.data_as(C.c_void_p)))

    def get_boundary_fluxes(self, boundary_index: int, num_edges: int, accumulated: bool = True) -> np.ndarray:
        out = np.zeros((3, num_edges))
        check(lib().rdyb_operator_get_boundary_fluxes(self._h, C.c_int32(boundary_index), C.c_int32(num_edges), int(accumulated),
                                                      out.ctypes.data_as(C.c_void_p)))
        return out

    def set_external_source(self, component: int, values):
        v, p = _f64(values)
        check(lib().rdyb_operator_set_external_source(self._h, C.c_int32(component), C.c_int32(v.shape[0]), p))

    def get_external_source(self, component: int) -> np.ndarray:
        out = np.zeros(self.num_owned)
        check(lib().rdyb_operator_get_external_source(self._h, C.c_int32(component), C.c_int32(self.num_owned), out.ctypes.data_as(C.c_void_p)))
        return out

    def set_regional_external_source(self, component: int, owned_cell_ids, values):
        ids = np.ascontiguousarray(owned_cell_ids, dtype=np.int32)
        v, p = _f64(values)
        check(lib().rdyb_operator_set_regional_external_source(self._h, C.c_int32(component), C.c_int32(len(ids)), ids.ctypes.data_as(C.c_void_p), p))

    def set_mannings(self, values):
        v, p = _f64(values)
        check(lib().rdyb_operator_set_mannings(self._h, C.c_int32(v.shape[0]), p))

    def set_regional_mannings(self, owned_cell_ids, values):
        ids = np.ascontiguousarray(owned_cell_ids, dtype=np.int32)
        v, p = _f64(values)
        check(lib().rdyb_operator_set_regional_mannings(self._h, C.c_int32(len(ids)), ids.ctypes.data_as(C.c_void_p), p))

    def set_state(self, u_global):
        v, p = _f64(u_global)
        assert v.size == 3 * self.num_owned, "u_global must hold 3*num_owned values"
        check(lib().rdyb_operator_set_state(self._h, p))

    def get_state(self) -> np.ndarray:
        out = np.zeros((self.num_owned, 3))
        check(lib().rdyb_operator_get_state(self._h, out.ctypes.data_as(C.c_void_p)))
        return out

    def get_state_component(self, component: int) -> np.ndarray:
        out = np.zeros(self.num_owned)
        check(lib().rdyb_operator_get_state_component(self._h, C.c_int32(component), C.c_int32(self.num_owned), out.ctypes.data_as(C.c_void_p)))
        return out

    def get_field(self, which: int) -> np.ndarray:
        out = np.zeros((self.num_owned, 3))
        check(lib().rdyb_operator_get_field(self._h, C.c_int32(which), out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- hot path ----
    def apply_host(self, dt: float, u_local: np.ndarray, f_global: np.ndarray | None = None) -> np.ndarray:
        """PetscOperator-apply shape with host Vec arrays: F = flux + sources of u_local."""
        u, pu = _f64(u_local)
        assert u.size == 3 * self.num_local
        if f_global is None:
            f_global = np.zeros((self.num_owned, 3))
        check(lib().rdyb_operator_apply_host(self._h, C.c_double(dt), pu, f_global.ctypes.data_as(C.c_void_p)))
        return f_global

    def apply_host_ptr(self, dt: float, u_ptr: int, f_ptr: int):
        check(lib().rdyb_operator_apply_host(self._h, C.c_double(dt), C.c_void_p(u_ptr), C.c_void_p(f_ptr)))

    def apply(self, dt: float, download: bool = True):
        if download:
            out = np.zeros((self.num_owned, 3))
            check(lib().rdyb_operator_apply(self._h, C.c_double(dt), out.ctypes.data_as(C.c_void_p)))
            return out
        check(lib().rdyb_operator_apply(self._h, C.c_double(dt), None))
        return None

    def step_euler(self, dt: float, nsteps: int = 1):
        check(lib().rdyb_operator_step_euler(self._h, C.c_double(dt), C.c_int32(nsteps)))

    def step_rk4(self, dt: float, nsteps: int = 1):
        check(lib().rdyb_operator_step_rk4(self._h, C.c_double(dt), C.c_int32(nsteps)))

    # ---- Bradford-Sanders well balancing (libCEED semantics) ----
    def enable_bs2002(self, mesh=None):
        """Vertex-eta pass + the g dhv^2 / 24 edge term (SWEEtaVertex / ComputeDhv of the reference's libCEED operator)."""
        mesh = mesh or self._keep[1]
        z = np.ascontiguousarray(mesh.vertices()[:, 2], dtype=np.float64)
        offs = np.ascontiguousarray(mesh.array("cell_offsets"), dtype=np.int32)
        ids = np.ascontiguousarray(mesh.array("cell_vertex_ids"), dtype=np.int32)
        ev = np.ascontiguousarray(mesh.array("edge_vertex_ids"), dtype=np.int32)
        i32 = C.POINTER(C.c_int32)
        vv = RdybVertexView(len(z), z.ctypes.data_as(C.POINTER(C.c_double)), offs.ctypes.data_as(i32), ids.ctypes.data_as(i32), ev.ctypes.data_as(i32))
        self._bs_keep = (z, offs, ids, ev)
        check(lib().rdyb_operator_enable_bs2002(self._h, C.byref(vv)))
        self.num_vertices = len(z)

    def eta_vertices(self) -> np.ndarray:
        out = np.zeros(self.num_vertices)
        check(lib().rdyb_operator_get_eta_vertices(self._h, C.c_int32(self.num_vertices), out.ctypes.data_as(C.c_void_p)))
        return out

    # ---- device-side time loop ----
    def time_set(self, t: float, dt: float):
        check(lib().rdyb_operator_time_set(self._h, C.c_double(t), C.c_double(dt)))

    def time_get(self):
        """(t, dt, steps, last Courant number) of the device-resident time control; waits for the queued intervals."""
        t, dt, c, n = C.c_double(), C.c_double(), C.c_double(), C.c_int64()
        check(lib().rdyb_operator_time_get(self._h, C.byref(t), C.byref(dt), C.byref(n), C.byref(c)))
        return t.value, dt.value, n.value, c.value

    def advance_device(self, interval: float, adaptive: bool = False, target_courant: float = 0.5, max_increase_factor: float = 1.1) -> bool:
        """One coupling interval as a single graph launch with a device-side loop; False when this configuration has none."""
        code = lib().rdyb_operator_advance_device(self._h, C.c_double(interval), int(adaptive), C.c_double(target_courant), C.c_double(max_increase_factor))
        if code == 56:  # RDYB_ERR_SUP
            return False
        check(code)
        return True

    def synchronize(self):
        check(lib().rdyb_operator_synchronize(self._h))

    def last_kernel_ms(self):
        ms, n = C.c_double(), C.c_int32()
        check(lib().rdyb_operator_last_kernel_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    # ---- diagnostics ----
    def kernel_info(self) -> str:
        buf = C.create_string_buffer(512)
        check(lib().rdyb_operator_kernel_info(self._h, buf, C.c_int32(512)))
        return buf.value.decode()

    def reset_diagnostics(self):
        check(lib().rdyb_operator_reset_diagnostics(self._h))

    def courant(self) -> RdybCourantDiagnostics:
        d = RdybCourantDiagnostics()
        check(lib().rdyb_operator_get_courant(self._h, C.byref(d)))
        return d

    def mass(self) -> float:
        m = C.c_double()
        check(lib().rdyb_operator_get_mass(self._h, C.byref(m)))
        return m.value

    # ---- multi-GPU ----
    @staticmethod
    def nccl_unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib().rdyb_nccl_unique_id(buf))
        return buf.raw

    def init_comm(self, nranks: int, rank: int, unique_id: bytes):
        check(lib().rdyb_operator_init_comm(self._h, C.c_int32(nranks), C.c_int32(rank), C.c_char_p(unique_id)))

    HALO_BLOB_BYTES = 512

    def halo_export(self, rank: int) -> bytes:
        buf = C.create_string_buffer(self.HALO_BLOB_BYTES)
        check(lib().rdyb_operator_halo_export(self._h, C.c_int32(rank), buf))
        return buf.raw

    def halo_import(self, nranks: int, rank: int, blobs: list) -> None:
        """blobs: the halo_export() result of every rank, in rank order (gathered by the caller, e.g. all_gather_object)."""
        raw = b"".join(blobs)
        assert len(raw) == nranks * self.HALO_BLOB_BYTES
        check(lib().rdyb_operator_halo_import(self._h, C.c_int32(nranks), C.c_int32(rank), raw))

    def halo_exchange(self):
        check(lib().rdyb_operator_halo_exchange(self._h))

    def close(self):
        if self._h:
            lib().rdyb_operator_destroy(C.byref(self._h))
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def host_sync_count() -> int:
    return int(lib().rdyb_host_sync_count())


def kernel_launch_count() -> int:
    return int(lib().rdyb_kernel_launch_count())
