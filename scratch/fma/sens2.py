"""Oracle-only: FMA-level perturbation growth, van Leer on the wet triangle floodplain (Euler vs RK4)."""
import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from oracle import binding
from rdycore_b200 import cases
from util import make_oracle, state_error
a_lib = binding.load("port")
orig = binding.PORT_LIB
binding.PORT_LIB = "/root/repo/scratch/fma/libswe_oracle_port.so"
binding._libs.pop("port"); f_lib = binding.load("port")
binding._libs["port"] = a_lib; binding.PORT_LIB = orig
def wet(recon):
    case = cases.floodplain(nx=40, ny=36, eta0=40.0)
    case.config.reconstruction = recon
    rng = np.random.default_rng(1)
    case.u0[:, 1] = 0.3 * rng.normal(size=len(case.u0)) * case.u0[:, 0]
    case.u0[:, 2] = 0.3 * rng.normal(size=len(case.u0)) * case.u0[:, 0]
    return case
for recon in (1, 3):
  for method, steps in (("euler", 60), ("euler", 100), ("euler",200), ("rk4", 40), ("rk4", 100)):
    case = wet(recon)
    o = make_oracle(case, "port")
    binding._libs["port"] = f_lib
    f = make_oracle(case, "port")
    binding._libs["port"] = a_lib
    for x in (o, f):
        (x.step_euler if method == "euler" else x.step_rk4)(case.dt, steps)
    if np.isnan(o.array("u_global")).any(): print(recon, method, steps, "nan"); continue
    e = state_error(f.array("u_global"), o.array("u_global"))
    print(recon, method, steps, "max_rel %.3e normwise %.3e" % (e[0], e[3]))
