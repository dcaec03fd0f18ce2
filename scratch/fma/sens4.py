"""Oracle-only: FMA-level perturbation growth on the reference decks."""
import sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
from oracle import binding
from util import make_oracle, state_error
import test_reference_decks as T
a_lib = binding.load("port")
orig = binding.PORT_LIB
binding.PORT_LIB = "/root/repo/scratch/fma/libswe_oracle_port.so"
binding._libs.pop("port"); f_lib = binding.load("port")
binding._libs["port"] = a_lib; binding.PORT_LIB = orig
for name in sorted(T.DECKS):
    case = T.deck_case(name)
    for steps in sorted({100, 300, T.DECKS[name]["steps"]}):
        if steps > T.DECKS[name]["steps"]: continue
        o = make_oracle(case, "port")
        binding._libs["port"] = f_lib
        f = make_oracle(case, "port")
        binding._libs["port"] = a_lib
        o.step_euler(case.dt, steps); f.step_euler(case.dt, steps)
        e = state_error(f.array("u_global"), o.array("u_global"))
        u = o.array("u_global").reshape(-1,3)
        print(name, steps, "max_rel %.3e normwise %.3e mismatch %d; hmin %.3g hmax %.3g qmax %.3g courant %.3f" % (e[0], e[3], e[2], u[:,0].min(), u[:,0].max(), np.abs(u[:,1:]).max(), o.courant()))
