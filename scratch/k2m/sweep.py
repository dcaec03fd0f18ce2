"""Block-size / residency / tile-size sweep of the tiled MUSCL kernel (K2M) on the second-order bench case (10 M cells, minmod).
Usage (GPU box): python scratch/k2m/sweep.py > gpurun_out/k2m_sweep.txt
The variants were selected by RDYB_K2M_VARIANT, a switch in tiled_muscl_launch that existed only for this sweep (results:
profiles/r1_k2m_sweep.txt); re-create it from the THREADS / MINB template parameters of swe_tiled_muscl_kernel to repeat."""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from rdycore_b200 import cases
from rdycore_b200.operator import GATHER, TILED, Operator

nx = int(os.environ.get("NX", 2236))
case = cases.floodplain(nx=nx, ny=nx, eta0=60.0)
case.config.reconstruction = cases.MUSCL_MINMOD
if os.environ.get("SWEEP", "1") == "1":
    configs = [("", GATHER, 224)] + [(v, TILED, t) for v, t in (("", 224), ("1", 224), ("1", 160), ("2", 128), ("2", 224), ("3", 128), ("3", 160), ("4", 128))]
else:
    configs = [(v, TILED, t) for v, t in (("4", 160), ("4", 192), ("5", 128), ("6", 128), ("7", 192), ("5", 160))]
for var, kind, tile in configs:
    if var:
        os.environ["RDYB_K2M_VARIANT"] = var
    else:
        os.environ.pop("RDYB_K2M_VARIANT", None)
    op = Operator.from_case(case)
    op.set_kernel(kind, tile)
    op.set_options(courant=True)
    op.step_euler(case.dt, 10)
    op.synchronize()
    t0 = time.perf_counter()
    op.step_euler(case.dt, 60)
    op.synchronize()
    ms = (time.perf_counter() - t0) / 60 * 1e3
    print(f"variant={var or 0} kind={kind} tile={tile}: {ms:.4f} ms/step, {op.num_owned / ms / 1e6:.2f} Gcell/s, mass {op.mass():.6e}; {op.kernel_info()[:60]}", flush=True)
    del op
