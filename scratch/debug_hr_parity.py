"""debug: where does the N-rank HR bench workload differ from the partitioned oracle?  torchrun --nproc-per-node 2 scratch/debug_hr_parity.py --nx 3536"""
import argparse, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=3536); ap.add_argument("--ny", type=int, default=0); ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--halo", default="p2p"); ap.add_argument("--kernel", type=int, default=3); ap.add_argument("--eta0", type=float, default=12.0)
ap.add_argument("--multi", action="store_true", help="bench flow: set_options, step_euler(dt, steps-1), reset_diagnostics, step_euler(dt, 1), one comparison")
ap.add_argument("--no-dam", action="store_true"); ap.add_argument("--no-rain", action="store_true"); ap.add_argument("--fo", action="store_true")
args = ap.parse_args()
import torch, torch.distributed as dist
from rdycore_b200 import cases
from rdycore_b200.operator import Operator
from oracle.partitioned import DistributedPart
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
px, py = {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}[world]
dt = 0.026798462824230505
ny = args.ny or args.nx
kw = dict(nx=args.nx, ny=ny, eta0=args.eta0, hr=not args.fo)
if not args.no_dam: kw["dam_column"] = (2000.0, 28.0)
case = cases.floodplain(px=px, py=py, rank=rank, dt=dt, **kw)
if args.no_rain: case.sources[:] = 0.0
op = Operator.from_case(case, device=local, halo=case.mesh.halo())
if args.kernel: op.set_kernel(args.kernel, 224)
uid = [Operator.nccl_unique_id() if rank == 0 else None]
dist.broadcast_object_list(uid, src=0)
op.init_comm(world, rank, uid[0])
if args.halo == "p2p":
    blobs = [None] * world
    dist.all_gather_object(blobs, op.halo_export(rank))
    op.halo_import(world, rank, blobs)
ref = DistributedPart(case, dist, rank)
m = case.mesh
o2l = m.array("owned_to_local")
gids = m.array("cell_global_ids")[o2l]
NX = args.nx
if args.multi:
    op.set_options(courant=True)
for s in range(1 if args.multi else args.steps):
    if args.multi:
        op.step_euler(dt, args.steps - 1); op.reset_diagnostics(); op.step_euler(dt, 1); op.synchronize()
        ref.step_euler(dt, args.steps)
    else:
        op.step_euler(dt, 1); op.synchronize()
        ref.step_euler(dt, 1)
    u, ur = op.get_state(), ref.u_global()
    d = np.abs(u - ur)
    worst = np.argsort(d.max(axis=1))[-3:][::-1]
    msg = f"[rank {rank}] step {s+1}: max|dh| {d[:,0].max():.3e} max|dq| {d[:,1:].max():.3e} n(|d|>1e-10) {(d.max(axis=1)>1e-10).sum()} kernel {op.kernel_info().split(';')[0]}"
    for w in worst:
        g = int(gids[w]); q = g // 2
        msg += f"\n    gid {g} quad (i={q % NX}, j={q // NX}) tri {g % 2}: gpu {u[w]} ref {ur[w]}"
    allm = [None] * world
    dist.all_gather_object(allm, msg)
    if rank == 0:
        print("\n".join(allm), flush=True)
op.close(); dist.barrier(); dist.destroy_process_group()
