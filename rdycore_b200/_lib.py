"""ctypes binding of librdyb200.so — the C-ABI declared in include/rdyb200.h and include/rdyb200_mesh.h.

There is no CPU fallback: if the shared library (with its CUDA kernels) is missing, loading fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RDYB_LIB") or os.path.join(HERE, "librdyb200.so")  # RDYB_LIB: a variant build, for experiments

c_i32p = C.POINTER(C.c_int32)
c_i64p = C.POINTER(C.c_int64)
c_f64p = C.POINTER(C.c_double)


class RdybConfig(C.Structure):
    _fields_ = [
        ("tiny_h", C.c_double),
        ("h_anuga_regular", C.c_double),
        ("xq2018_threshold", C.c_double),
        ("source_method", C.c_int32),
        ("well_balancing", C.c_int32),
        ("temporal", C.c_int32),
        ("reconstruction", C.c_int32),
    ]


class RdybMeshView(C.Structure):
    _fields_ = [
        ("num_cells", C.c_int32),
        ("num_owned_cells", C.c_int32),
        ("num_edges", C.c_int32),
        ("num_internal_edges", C.c_int32),
        ("cell_is_owned", c_i32p),
        ("cell_local_to_owned", c_i32p),
        ("cell_global_ids", c_i64p),
        ("cell_areas", c_f64p),
        ("cell_dz_dx", c_f64p),
        ("cell_dz_dy", c_f64p),
        ("cell_zc", c_f64p),
        ("cell_xc", c_f64p),
        ("cell_yc", c_f64p),
        ("edge_cell_ids", c_i32p),
        ("edge_lengths", c_f64p),
        ("edge_cn", c_f64p),
        ("edge_sn", c_f64p),
        ("internal_edge_ids", c_i32p),
        ("edge_global_ids", c_i64p),
        ("edge_xm", c_f64p),
        ("edge_ym", c_f64p),
    ]


class RdybBoundaryView(C.Structure):
    _fields_ = [
        ("id", C.c_int32),
        ("index", C.c_int32),
        ("num_edges", C.c_int32),
        ("edge_ids", c_i32p),
        ("bc_type", C.c_int32),
    ]


class RdybCourantDiagnostics(C.Structure):
    _fields_ = [("max_courant_num", C.c_double), ("global_edge_id", C.c_int64), ("global_cell_id", C.c_int64)]


class RdybVertexView(C.Structure):
    _fields_ = [("num_vertices", C.c_int32), ("vertex_z", C.POINTER(C.c_double)), ("cell_vertex_offsets", c_i32p), ("cell_vertex_ids", c_i32p),
                ("edge_vertex_ids", c_i32p)]


class RdybHaloView(C.Structure):
    _fields_ = [
        ("num_peers", C.c_int32),
        ("peer_ranks", c_i32p),
        ("send_offsets", c_i32p),
        ("send_cells", c_i32p),
        ("recv_offsets", c_i32p),
        ("recv_cells", c_i32p),
    ]


# every symbol include/*.h declares; tests/test_abi.py checks that the library exports each one
MESH_SYMBOLS = [
    "rdyb_hostmesh_create", "rdyb_hostmesh_destroy", "rdyb_hostmesh_override_2d_projection", "rdyb_hostmesh_set_zc",
    "rdyb_hostmesh_view", "rdyb_hostmesh_halo", "rdyb_hostmesh_sizes", "rdyb_hostmesh_array", "rdyb_hostmesh_partition_rcb",
    "rdyb_hostmesh_extract_local", "rdyb_hostmesh_generate_quads", "rdyb_hostmesh_generate_floodplain",
    "rdyb_hostmesh_refine_triangles", "rdyb_hostmesh_set_vertex_z", "rdyb_hostmesh_vertices",
]
OPERATOR_SYMBOLS = [
    "rdyb_last_error", "rdyb_version", "rdyb_kernel_launch_count", "rdyb_operator_create", "rdyb_operator_destroy",
    "rdyb_operator_num_owned_cells", "rdyb_operator_num_local_cells", "rdyb_operator_set_options",
    "rdyb_operator_set_boundary_values", "rdyb_operator_get_boundary_fluxes", "rdyb_operator_set_external_source",
    "rdyb_operator_get_external_source", "rdyb_operator_set_regional_external_source", "rdyb_operator_set_mannings",
    "rdyb_operator_set_regional_mannings", "rdyb_operator_set_state", "rdyb_operator_get_state",
    "rdyb_operator_get_state_component", "rdyb_operator_apply_host", "rdyb_operator_apply", "rdyb_operator_step_euler",
    "rdyb_operator_step_rk4", "rdyb_operator_synchronize", "rdyb_operator_reset_diagnostics", "rdyb_operator_get_courant",
    "rdyb_operator_get_mass", "rdyb_nccl_unique_id", "rdyb_operator_init_comm", "rdyb_operator_halo_exchange",
    "rdyb_operator_device_pointer", "rdyb_operator_last_kernel_ms", "rdyb_operator_set_kernel", "rdyb_operator_get_field",
    "rdyb_host_alloc", "rdyb_host_free", "rdyb_operator_kernel_info", "rdyb_operator_halo_export", "rdyb_operator_halo_import",
    "rdyb_operator_set_external_source_aos", "rdyb_operator_get_courant_global", "rdyb_operator_set_max_ctas",
    "rdyb_operator_get_wait_cycles", "rdyb_operator_time_set", "rdyb_operator_time_get", "rdyb_operator_advance_device", "rdyb_host_sync_count", "rdyb_operator_enable_bs2002", "rdyb_operator_get_eta_vertices",
    "rdyb_mass_sum_noalloc",
]

_lib = None


class RdybError(RuntimeError):
    """Raised when a C-ABI call returns a non-zero code (PetscErrorCode-like)."""

    def __init__(self, code: int, message: str):
        super().__init__(f"[rdyb error {code}] {message}")
        self.code = code


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m rdycore_b200.build` "
                "(there is no CPU fallback for the SWE hot path)"
            )
        _lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
        _lib.rdyb_last_error.restype = C.c_char_p
        if hasattr(_lib, "rdyb_kernel_launch_count"):
            _lib.rdyb_kernel_launch_count.restype = C.c_int64
        if hasattr(_lib, "rdyb_host_sync_count"):
            _lib.rdyb_host_sync_count.restype = C.c_int64
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise RdybError(code, lib().rdyb_last_error().decode(errors="replace"))
