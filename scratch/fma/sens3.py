"""Oracle-only: FMA-level perturbation growth on the second-order MMS problem."""
import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from oracle import binding
from rdycore_b200 import cases
from util import make_oracle, state_error
import test_mms as T
a_lib = binding.load("port")
orig = binding.PORT_LIB
binding.PORT_LIB = "/root/repo/scratch/fma/libswe_oracle_port.so"
binding._libs.pop("port"); f_lib = binding.load("port")
binding._libs["port"] = a_lib; binding.PORT_LIB = orig
def run(case, lib, nsteps):
    binding._libs["port"] = lib
    o = make_oracle(case, "port")
    binding._libs["port"] = a_lib
    i = case.info
    for k in range(nsteps):
        t = k * T.DT + 0.5 * T.DT
        o.array("boundary_values", 0)[:] = T.FORMS.solution(i["xe"], i["ye"], t).ravel()
        o.array("external_sources")[:] = T.FORMS.sources(i["x"], i["y"], t).ravel()
        o.step_euler(T.DT, 1)
    return o.array("u_global").reshape(-1, 3).copy()
for recon in (0, 1):
  for level in (1, 2, 3):
    for n in (100, 500):
      case = T.mms_case(level, recon)
      e = state_error(run(case, f_lib, n), run(case, a_lib, n))
      print("recon", recon, "level", level, "steps", n, "max_rel %.3e normwise %.3e" % (e[0], e[3]))
