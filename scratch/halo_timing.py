"""torchrun worker: time the pieces of the N>1 step (halo exchange alone, step without comm, full step)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from rdycore_b200 import cases
from rdycore_b200.operator import Operator
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
px, py = {2: (2, 1), 4: (2, 2), 8: (4, 2)}[world]
nx = int(os.environ.get("NX", "2236"))
case = cases.floodplain(nx=nx * px, ny=nx * py, eta0=18.0, px=px, py=py, rank=rank)
op = Operator.from_case(case, device=local, halo=case.mesh.halo())
op.set_kernel(3, 224)
def timeit(fn, n):
    op.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter(); fn(n); op.synchronize(); torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3
# without a communicator: interior + frontier launches only (ghosts stale)
op.step_euler(case.dt, 5)
t_nocomm = timeit(lambda n: op.step_euler(case.dt, n), 200)
if rank == 0: print("HALO_TIMING", op.kernel_info(), flush=True)
uid = [Operator.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
op.init_comm(world, rank, uid[0])
op.step_euler(case.dt, 5)
t_full = timeit(lambda n: op.step_euler(case.dt, n), 200)
def halo(n):
    for _ in range(n): op.halo_exchange()
halo(5)
t_halo = timeit(halo, 200)
blobs = [None] * world
dist.all_gather_object(blobs, op.halo_export(rank))
op.halo_import(world, rank, blobs)
op.step_euler(case.dt, 5)
t_p2p = timeit(lambda n: op.step_euler(case.dt, n), 200)
if rank == 0:
    print(f"HALO_TIMING p2p full step {t_p2p:.4f} ms", flush=True)
    print(f"HALO_TIMING ranks={world} owned={op.num_owned} local={op.num_local} ms/step: no-comm {t_nocomm:.4f}  full {t_full:.4f}  halo-exchange alone {t_halo:.4f}", flush=True)
op.close(); dist.barrier(); dist.destroy_process_group()
