"""CPU: the C-ABI library loads and exports every symbol include/*.h declares; error paths behave like PetscCheck."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from rdycore_b200 import _lib, cases
from rdycore_b200.mesh import HostMesh

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b((?:rdyb_|RDy[A-Z])\w*)\s*\(", txt)))


@pytest.mark.parametrize("header", [h for h in sorted(os.listdir(os.path.join(ROOT, "include"))) if h.endswith(".h")])
def test_every_declared_symbol_is_exported(header):
    lib = _lib.lib()
    names = _declared(header)
    assert names, f"no declarations parsed from {header}"
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"{header}: librdyb200.so does not export {missing}"


def test_python_symbol_lists_match_headers():
    declared = set(_declared("rdyb200.h")) | set(_declared("rdyb200_mesh.h"))
    listed = set(_lib.MESH_SYMBOLS) | set(_lib.OPERATOR_SYMBOLS)
    assert listed <= declared, f"listed but not declared: {sorted(listed - declared)}"


def test_version_and_error_text():
    lib = _lib.lib()
    major, minor = C.c_int(), C.c_int()
    assert lib.rdyb_version(C.byref(major), C.byref(minor)) == 0
    # a bad cell (5 vertices) is rejected with the reference's message (src/rdymesh.c:811)
    xyz = np.zeros((5, 3))
    with pytest.raises(_lib.RdybError) as ei:
        HostMesh.from_cells(xyz, np.array([0, 5], np.int32), np.arange(5, dtype=np.int32))
    assert ei.value.code == 83 and "must be 3 or 4" in str(ei.value)


def test_operator_create_fails_loudly_without_a_gpu():
    """No CPU fallback: without a CUDA device the operator refuses to exist."""
    try:
        import torch

        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    case = cases.dam_break_quads(8, 8)
    from rdycore_b200.operator import Operator

    with pytest.raises(_lib.RdybError) as ei:
        Operator.from_case(case)
    assert ei.value.code == 97 and "no CPU fallback" in str(ei.value)


def test_invalid_configuration_is_rejected_before_touching_the_gpu():
    case = cases.dam_break_quads(8, 8)
    case.config.source_method = 2  # ark_imex: "Only semi_implicit and implicit_xq2018 are supported in the PETSc version"
    from rdycore_b200.operator import Operator

    with pytest.raises(_lib.RdybError) as ei:
        Operator.from_case(case)
    assert ei.value.code == 83 and "semi_implicit" in str(ei.value)
