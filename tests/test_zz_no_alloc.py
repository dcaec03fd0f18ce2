"""rdyb_mass_sum_noalloc: the allocation-free variant of rdyb_operator_get_mass (include/rdyb200.h, csrc/swe_noalloc.cu).

No counterpart in the reference.  The file name sorts last on purpose: the entry point
shares one static device array between all calls of the process, so it is exercised after everything else and must pass both alone in a
fresh process and at the end of the whole suite.
"""
import ctypes as C
import os
import re

import numpy as np
import pytest

from rdycore_b200 import _lib, cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MAX_CELLS = 1 << 30  # RDYB_NOALLOC_MAX_CELLS
REL_TOL = 1e-13  # the summation ORDER differs (owned order here, the operator's device order there): a few ulp of a 1000-term sum


def _entry():
    f = _lib.lib().rdyb_mass_sum_noalloc
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    f.restype = C.c_int
    return f


def test_declared_limit_and_source_allocate_nothing():
    hdr = open(os.path.join(ROOT, "include", "rdyb200.h")).read()
    assert int(re.search(r"#define RDYB_NOALLOC_MAX_CELLS (\d+)", hdr).group(1)) == MAX_CELLS
    src = open(os.path.join(ROOT, "rdycore_b200", "csrc", "swe_noalloc.cu")).read()
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMemcpy", "Synchronize", "malloc(", "new "):
        assert banned not in code, f"swe_noalloc.cu calls {banned}"
    assert "static __device__ __align__(256)" in code and "cudaGetSymbolAddress" in code


def test_sizes_beyond_the_maximum_are_rejected_before_any_cuda_call():
    f = _entry()
    out = C.c_double(0.0)  # never dereferenced: the size check comes first
    assert f(None, None, C.c_int64(MAX_CELLS + 1), C.byref(out), None) == 60  # RDYB_ERR_ARG_SIZ
    msg = _lib.lib().rdyb_last_error().decode()
    assert "exceed the maximum" in msg and str(MAX_CELLS) in msg
    assert f(None, None, C.c_int64(-1), C.byref(out), None) == 60
    assert f(None, None, C.c_int64(16), None, None) == 62  # RDYB_ERR_ARG: NULL output


@pytest.mark.gpu
def test_noalloc_mass_equals_the_operator_mass_on_two_calls_in_a_row():
    import torch

    from rdycore_b200.operator import TILED_FIXED, Operator

    f = _entry()
    case = cases.floodplain(nx=96, ny=80, eta0=8.0)
    mesh = case.mesh
    area_owned = mesh.view_arrays()["cell_areas"][mesh.array("owned_to_local")]
    op = Operator.from_case(case, device=0)
    op.set_kernel(TILED_FIXED, 224)
    torch.cuda.set_device(0)
    area = torch.from_numpy(np.ascontiguousarray(area_owned)).cuda()
    h = torch.empty(op.num_owned, dtype=torch.float64, device="cuda")
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    stream = torch.cuda.current_stream()
    results = []
    try:
        for call in range(2):  # two calls in a row on different states: the second one meets the first one's partial sums in the scratch
            op.step_euler(case.dt, 40)
            m_ref = op.mass()  # the existing entry point (cudaMalloc'ed scratch, host-side final sum)
            u = op.get_state()
            h.copy_(torch.from_numpy(np.ascontiguousarray(u[:, 0])))
            out.fill_(float("nan"))
            ierr = f(h.data_ptr(), area.data_ptr(), C.c_int64(op.num_owned), out.data_ptr(), C.c_void_p(stream.cuda_stream))
            assert ierr == 0, _lib.lib().rdyb_last_error().decode()
            stream.synchronize()
            m = float(out.item())
            assert np.isfinite(m) and m_ref > 0.0
            assert abs(m - m_ref) <= REL_TOL * m_ref, (call, m, m_ref)
            assert abs(m - float(np.dot(area_owned, u[:, 0]))) <= 1e-12 * m_ref
            results.append(m)
        assert results[0] != results[1]  # rain fell in between: the two calls really summed different states
        # a shorter input after a longer one: no partial sum of the previous call may leak into this one; repeated -> bitwise equal
        n_small = 1000
        expect = float(np.dot(area_owned[:n_small], u[:n_small, 0]))
        small = []
        for _ in range(2):
            out.fill_(float("nan"))
            assert f(h.data_ptr(), area.data_ptr(), C.c_int64(n_small), out.data_ptr(), C.c_void_p(stream.cuda_stream)) == 0
            stream.synchronize()
            small.append(float(out.item()))
        assert small[0] == small[1] and abs(small[0] - expect) <= REL_TOL * abs(expect)
        # an empty input sums to exactly zero, and an oversized one launches nothing and leaves the output alone
        assert f(None, None, C.c_int64(0), out.data_ptr(), C.c_void_p(stream.cuda_stream)) == 0
        stream.synchronize()
        assert float(out.item()) == 0.0
        out.fill_(-1.0)
        assert f(h.data_ptr(), area.data_ptr(), C.c_int64(MAX_CELLS + 1), out.data_ptr(), C.c_void_p(stream.cuda_stream)) == 60
        stream.synchronize()
        assert float(out.item()) == -1.0
    finally:
        op.close()
