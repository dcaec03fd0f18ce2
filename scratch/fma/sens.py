"""Oracle-only: how far does an FMA-level rounding perturbation grow on the second-order dam break?"""
import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from oracle import binding
from rdycore_b200 import cases
from util import make_oracle, state_error
binding.PORT_LIB_FMA = "/root/repo/scratch/fma/libswe_oracle_port.so"
a_lib = binding.load("port")
import ctypes as C
# load FMA variant under a fake kind
orig = binding.PORT_LIB
binding.PORT_LIB = binding.PORT_LIB_FMA
binding._libs.pop("port")
f_lib = binding.load("port")
binding._libs["port"] = a_lib; binding.PORT_LIB = orig
for recon in (1, 2, 3):
    for steps in (100, 200, 400):
        case = cases.dam_break_quads(nx=60, ny=60)
        case.config.reconstruction = recon
        o = make_oracle(case, "port")
        binding._libs["port"] = f_lib
        f = make_oracle(case, "port")
        binding._libs["port"] = a_lib
        o.step_euler(case.dt, steps); f.step_euler(case.dt, steps)
        print(recon, steps, "max_rel %.3e normwise %.3e" % (state_error(f.array("u_global"), o.array("u_global"))[0], state_error(f.array("u_global"), o.array("u_global"))[3]))
