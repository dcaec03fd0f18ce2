"""ctypes mirror of include/rdyb200_petsc.h: the PetscOperator sub-operator interface and the B200 flux / source factories.

Names follow the reference (src/petsc_operator.c, include/private/rdyoperatorimpl.h:45-90) so that tests read like
RDycore code: PetscOperatorCreate, PetscOperatorCompositeAddSub, PetscOperatorSetField, PetscOperatorApply ...
Everything computes through the C-ABI; nothing here does arithmetic.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import RdybCourantDiagnostics, check, lib

MAX_NAME_LEN = 128


class Vec:
    """Host Vec stand-in (array + local size + block size), rdyb_vec_* in the library."""

    def __init__(self, local_size: int, block_size: int = 1, pinned: bool = False):
        self._h = C.c_void_p()
        check(lib().rdyb_vec_create(C.c_int64(local_size), C.c_int32(block_size), int(pinned), C.byref(self._h)))
        self.local_size, self.block_size = local_size, block_size

    @property
    def handle(self):
        return self._h

    def array(self) -> np.ndarray:
        """A numpy view of the Vec's storage, taken through VecGetArray/VecRestoreArray (so the state counter moves)."""
        p = C.POINTER(C.c_double)()
        check(lib().rdyb_vec_get_array(self._h, C.byref(p)))
        a = np.ctypeslib.as_array(p, shape=(max(self.local_size, 1),))[: self.local_size]
        check(lib().rdyb_vec_restore_array(self._h, C.byref(p)))
        return a

    def set(self, values):
        """VecGetArray; copy; VecRestoreArray."""
        p = C.POINTER(C.c_double)()
        check(lib().rdyb_vec_get_array(self._h, C.byref(p)))
        np.ctypeslib.as_array(p, shape=(max(self.local_size, 1),))[: self.local_size] = np.asarray(values, dtype=np.float64).ravel()
        check(lib().rdyb_vec_restore_array(self._h, C.byref(p)))

    def state(self) -> int:
        s = C.c_int64()
        check(lib().rdyb_vec_get_state(self._h, C.byref(s)))
        return s.value

    def zero(self):
        check(lib().rdyb_vec_zero(self._h))

    def copy_to(self, other: "Vec"):
        check(lib().rdyb_vec_copy(self._h, other._h))

    def destroy(self):
        if self._h:
            check(lib().rdyb_vec_destroy(C.byref(self._h)))
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class PetscOperatorField(C.Structure):
    _fields_ = [("name", C.c_char * (MAX_NAME_LEN + 1)), ("vec", C.c_void_p)]


class PetscOperatorFields(C.Structure):
    _fields_ = [("num_fields", C.c_int32), ("capacity", C.c_int32), ("fields", C.POINTER(PetscOperatorField))]

    def get(self, name: str):
        v = C.c_void_p()
        check(lib().rdyb_petscop_fields_get(self, name.encode(), C.byref(v)))
        return v.value


APPLY_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, PetscOperatorFields, C.c_double, C.c_void_p, C.c_void_p)
DESTROY_FN = C.CFUNCTYPE(C.c_int, C.c_void_p)


class _OperatorStruct(C.Structure):
    _fields_ = [("context", C.c_void_p), ("fields", PetscOperatorFields), ("is_composite", C.c_int), ("apply", APPLY_FN), ("destroy", DESTROY_FN)]


class PetscOperator:
    """Handle of a struct RdybPetscOperator_ (same layout as struct _p_PetscOperator)."""

    def __init__(self, handle: C.c_void_p, keep=()):
        self._h = handle
        self._keep = list(keep)
        self.owned_by_composite = False

    @property
    def struct(self) -> _OperatorStruct:
        return C.cast(self._h, C.POINTER(_OperatorStruct)).contents

    @property
    def is_composite(self) -> bool:
        return bool(self.struct.is_composite)


def _vec_handle(v):
    return v.handle if isinstance(v, Vec) else v


def PetscOperatorCreate(apply, destroy, context=None) -> PetscOperator:
    """apply(context, fields, dt, u_local_handle, f_global_handle) -> int; destroy(context) -> int (Python callables)."""
    a, d = APPLY_FN(apply), DESTROY_FN(destroy)
    h = C.c_void_p()
    check(lib().rdyb_petscop_create(C.c_void_p(context), a, d, C.byref(h)))
    return PetscOperator(h, keep=(a, d))


def PetscOperatorCreateComposite() -> PetscOperator:
    h = C.c_void_p()
    check(lib().rdyb_petscop_create_composite(C.byref(h)))
    return PetscOperator(h)


def PetscOperatorCompositeAddSub(op: PetscOperator, sub: PetscOperator):
    check(lib().rdyb_petscop_composite_add_sub(op._h, sub._h))
    sub.owned_by_composite = True
    op._keep.append(sub)


def PetscOperatorSetField(op: PetscOperator, name: str, vec):
    check(lib().rdyb_petscop_set_field(op._h, name.encode(), _vec_handle(vec)))
    op._keep.append(vec)


def PetscOperatorApply(op: PetscOperator, dt: float, u_local, f_global):
    check(lib().rdyb_petscop_apply(op._h, C.c_double(dt), _vec_handle(u_local), _vec_handle(f_global)))


def PetscOperatorDestroy(op: PetscOperator):
    if op._h and not op.owned_by_composite:
        check(lib().rdyb_petscop_destroy(C.byref(op._h)))
    op._h = C.c_void_p()


def CreateB200FluxOperator(config, mesh_view, boundaries, num_boundaries, boundary_values=None, boundary_fluxes=None, boundary_fluxes_accum=None,
                           diagnostics: RdybCourantDiagnostics | None = None, device: int = 0) -> PetscOperator:
    """rdyb_create_flux_operator: the replacement of CreatePetscFluxOperator / CreatePetscFluxHROperator
    (include/private/rdyoperatorimpl.h:234-237).  boundary_*: lists of Vec, one per boundary."""

    def vec_array(vs):
        if vs is None:
            return None
        arr = (C.c_void_p * max(len(vs), 1))(*[_vec_handle(v) for v in vs])
        return arr

    bv, bf, ba = vec_array(boundary_values), vec_array(boundary_fluxes), vec_array(boundary_fluxes_accum)
    h = C.c_void_p()
    check(lib().rdyb_create_flux_operator(C.byref(config), C.byref(mesh_view), C.c_int32(num_boundaries), boundaries, bv, bf, ba,
                                          C.byref(diagnostics) if diagnostics is not None else None, C.c_int(device), C.byref(h)))
    return PetscOperator(h, keep=(config, mesh_view, boundaries, bv, bf, ba, boundary_values, boundary_fluxes, boundary_fluxes_accum, diagnostics))


def CreateB200SourceOperator(config, mesh_view, external_sources: Vec, material_properties: Vec, flux_op: PetscOperator) -> PetscOperator:
    """rdyb_create_source_operator: the replacement of CreatePetscSourceOperator / CreatePetscSourceHROperator (:238-239)."""
    h = C.c_void_p()
    check(lib().rdyb_create_source_operator(C.byref(config), C.byref(mesh_view), external_sources.handle, material_properties.handle, flux_op._h,
                                            C.byref(h)))
    return PetscOperator(h, keep=(external_sources, material_properties, flux_op))


def set_fused(flux_op: PetscOperator, fused: bool):
    check(lib().rdyb_petscop_set_fused(flux_op._h, int(fused)))


def device_operator_handle(op: PetscOperator) -> C.c_void_p:
    h = C.c_void_p()
    check(lib().rdyb_petscop_get_device_operator(op._h, C.byref(h)))
    return h
