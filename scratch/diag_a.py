import sys, os
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
from rdycore_b200 import cases
from rdycore_b200.operator import Operator
from util import make_oracle, state_error
case = cases.dam_break_quads(100, 100, source_method=int(os.environ.get("SRC","1")))
oracle = make_oracle(case); op = Operator.from_case(case)
done=0
for n in (1,9,90,400,500):
    op.step_euler(case.dt, n); oracle.step_euler(case.dt, n); done+=n
    u=op.get_state(); r=oracle.array("u_global").reshape(-1,3)
    d=np.abs(u-r)
    out=[]
    for c in range(3):
        i=int(np.argmax(d[:,c])); out.append("c%d: maxabs %.2e at cell %d (ref %.3e) max|x| %.2e relinf %.2e"%(c,d[i,c],i,r[i,c],np.abs(r[:,c]).max(), d[:,c].max()/np.abs(r[:,c]).max()))
    print(done, state_error(u,r), " | ".join(out))
