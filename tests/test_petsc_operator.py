"""CPU: the PetscOperator object of include/rdyb200_petsc.h behaves like the reference's (src/petsc_operator.c).

No GPU compute here: sub-operators are Python callbacks.  The GPU-backed flux / source sub-operators are exercised
against the oracle in tests/test_gpu_petsc_operator.py.
"""
import ctypes as C

import numpy as np
import pytest

from rdycore_b200 import _lib, cases
from rdycore_b200 import petsc_operator as po


def _recording_op(log, tag, rc=0):
    def apply(ctx, fields, dt, u, f):
        log.append(("apply", tag, dt, u, f, fields.num_fields, [fields.fields[i].name.decode() for i in range(fields.num_fields)]))
        return rc

    def destroy(ctx):
        log.append(("destroy", tag))
        return 0

    return po.PetscOperatorCreate(apply, destroy)


def test_single_operator_apply_and_destroy():
    log = []
    op = _recording_op(log, "a")
    assert not op.is_composite
    u, f = po.Vec(6, 3), po.Vec(6, 3)
    po.PetscOperatorApply(op, 0.25, u, f)
    assert log[0][:3] == ("apply", "a", 0.25) and log[0][3] == u.handle.value and log[0][4] == f.handle.value
    po.PetscOperatorDestroy(op)
    assert log[-1] == ("destroy", "a")


def test_composite_applies_subs_in_insertion_order_and_owns_them():
    # PetscOperatorCompositeApply loops in insertion order (src/petsc_operator.c:54-60); the composite destroys its subs (:62-71)
    log = []
    comp = po.PetscOperatorCreateComposite()
    assert comp.is_composite
    for tag in "abcdefghij":  # more than the initial capacity of 8 (:89-92)
        po.PetscOperatorCompositeAddSub(comp, _recording_op(log, tag))
    u, f = po.Vec(3, 3), po.Vec(3, 3)
    po.PetscOperatorApply(comp, 1.5, u, f)
    assert [e[1] for e in log] == list("abcdefghij")
    log.clear()
    po.PetscOperatorDestroy(comp)
    assert sorted(e[1] for e in log if e[0] == "destroy") == list("abcdefghij")


def test_apply_error_propagates_and_stops_the_composite():
    log = []
    comp = po.PetscOperatorCreateComposite()
    po.PetscOperatorCompositeAddSub(comp, _recording_op(log, "ok"))
    po.PetscOperatorCompositeAddSub(comp, _recording_op(log, "bad", rc=83))
    po.PetscOperatorCompositeAddSub(comp, _recording_op(log, "never"))
    u, f = po.Vec(3, 3), po.Vec(3, 3)
    with pytest.raises(_lib.RdybError) as ei:
        po.PetscOperatorApply(comp, 1.0, u, f)
    assert ei.value.code == 83
    assert [e[1] for e in log] == ["ok", "bad"]
    po.PetscOperatorDestroy(comp)


def test_set_field_propagates_to_subs_and_replaces_by_name():
    # src/petsc_operator.c:130-166: set on the composite and on every sub; same name replaces; FieldsGet gives NULL when absent
    log = []
    comp = po.PetscOperatorCreateComposite()
    subs = [_recording_op(log, t) for t in "ab"]
    for s in subs:
        po.PetscOperatorCompositeAddSub(comp, s)
    v1, v2, v3 = po.Vec(3, 3), po.Vec(3, 3), po.Vec(3, 3)
    po.PetscOperatorSetField(comp, "riemannf", v1)
    po.PetscOperatorSetField(comp, "primitive_variables", v2)
    for o in [comp] + subs:
        assert o.struct.fields.num_fields == 2
        assert o.struct.fields.get("riemannf") == v1.handle.value
        assert o.struct.fields.get("primitive_variables") == v2.handle.value
        assert o.struct.fields.get("no_such_field") is None
    po.PetscOperatorSetField(comp, "riemannf", v3)
    for o in [comp] + subs:
        assert o.struct.fields.num_fields == 2 and o.struct.fields.get("riemannf") == v3.handle.value
    u, f = po.Vec(3, 3), po.Vec(3, 3)
    po.PetscOperatorApply(comp, 1.0, u, f)
    assert log[0][5] == 2 and log[0][6] == ["riemannf", "primitive_variables"]
    # a sub added later does not inherit earlier fields (the reference sets fields after all subs exist, src/operator.c:224-299)
    late = _recording_op(log, "late")
    po.PetscOperatorCompositeAddSub(comp, late)
    assert late.struct.fields.num_fields == 0
    po.PetscOperatorDestroy(comp)
    # the Vecs are not owned by the operator: still usable
    v1.set([1.0, 2.0, 3.0])
    assert v1.array().tolist() == [1.0, 2.0, 3.0]


def test_many_fields_grow_the_table():
    op = _recording_op([], "x")
    vecs = [po.Vec(1, 1) for _ in range(20)]
    for i, v in enumerate(vecs):
        po.PetscOperatorSetField(op, f"field_{i}", v)
    assert op.struct.fields.num_fields == 20
    assert op.struct.fields.get("field_17") == vecs[17].handle.value
    po.PetscOperatorDestroy(op)


def test_add_sub_to_a_non_composite_is_an_error():
    a, b = _recording_op([], "a"), _recording_op([], "b")
    with pytest.raises(_lib.RdybError):
        po.PetscOperatorCompositeAddSub(a, b)
    po.PetscOperatorDestroy(a)
    po.PetscOperatorDestroy(b)


def test_host_vec_semantics():
    v = po.Vec(12, 3)
    assert v.array().shape == (12,) and not v.array().any()
    s0 = v.state()
    v.set(np.arange(12.0))
    assert v.state() > s0  # VecRestoreArray bumps the object state: how the B200 operators notice changed inputs
    w = po.Vec(12, 3)
    v.copy_to(w)
    assert w.array().tolist() == list(np.arange(12.0))
    w.zero()
    assert not w.array().any()
    with pytest.raises(_lib.RdybError) as ei:
        po.Vec(10, 3)  # local size must be a multiple of the block size
    assert ei.value.code == 60
    with pytest.raises(_lib.RdybError):
        v.copy_to(po.Vec(9, 3))
    p = C.POINTER(C.c_double)()
    assert _lib.lib().rdyb_vec_restore_array(v.handle, C.byref(p)) == 83  # restore without get


def test_flux_factory_fails_loudly_without_a_gpu_and_source_needs_a_b200_flux_operator():
    import torch

    case = cases.dam_break_quads(nx=8, ny=8)
    arr, nb = case.boundary_views()
    view = case.mesh.view()
    if not torch.cuda.is_available():
        with pytest.raises(_lib.RdybError) as ei:
            po.CreateB200FluxOperator(case.config, view, arr, nb)
        assert ei.value.code == 97 and "no CPU fallback" in str(ei.value)
    n = view.num_owned_cells
    ext, mat = po.Vec(3 * n, 3), po.Vec(n, 1)
    not_b200 = _recording_op([], "cpu")
    with pytest.raises(_lib.RdybError) as ei:
        po.CreateB200SourceOperator(case.config, view, ext, mat, not_b200)
    assert ei.value.code == 83 and "share one GPU operator" in str(ei.value)
    po.PetscOperatorDestroy(not_b200)
