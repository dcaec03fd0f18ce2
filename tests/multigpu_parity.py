"""Multi-rank parity worker (launched by torchrun from tests/test_multi_gpu.py or by hand):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
      tests/multigpu_parity.py [--hr] [--steps 60] [--nx 96] [--ny 64] [--kernel 3] [--tile 224]

Every rank builds ITS block of the synthetic floodplain (one layer of face-neighbour ghosts, DMPlexDistributeOverlap(dm,1)
in the reference, src/rdydm.c:145-157), steps it on its own GPU with the per-step NCCL halo exchange
(DMGlobalToLocal, src/rdysetup.c:1133-1134) and the result is compared, by global cell id, with the CPU oracle run
serially on the whole mesh.  Tolerance as in tests/util.py (1e-11 relative, north_star)."""
from __future__ import annotations

import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--hr", action="store_true")
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--nx", type=int, default=96)
    ap.add_argument("--ny", type=int, default=64)
    ap.add_argument("--kernel", type=int, default=3)
    ap.add_argument("--tile", type=int, default=224)
    ap.add_argument("--p2p", action="store_true", help="peer-memory halo (CUDA IPC over NVLink) instead of NCCL send/recv")
    ap.add_argument("--rk4", action="store_true", help="classical RK4 instead of forward Euler: with --p2p the four stage vectors of a step travel over the "
                    "peer-memory halo (rdyb_operator_step_rk4: stage buffers alternate so that consecutive RHS evaluations never share one)")
    ap.add_argument("--mixed-paths", action="store_true", help="momentum sources on rank 0's cells only: rank 0 leaves the fused-push K4 path "
                    "(general kernel + push kernel) while the other ranks stay on it; the epoch accounting must keep them in step")
    ap.add_argument("--max-ctas", type=int, default=0, help="cap the persistent grid: every CTA walks many tiles (frontier and interior), the bench regime")
    ap.add_argument("--recon", type=int, default=0, help="second order: 1 minmod, 2 no limiter, 3 van Leer (fully wet case; gradients are exchanged too)")
    args = ap.parse_args()

    import torch
    import torch.distributed as dist

    from rdycore_b200 import cases
    from rdycore_b200.operator import Operator

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    px, py = {1: (1, 1), 2: (2, 1), 4: (2, 2), 8: (4, 2)}[world]
    kw = dict(nx=args.nx, ny=args.ny, eta0=40.0 if args.recon else 6.0 if args.hr else 8.0, hr=args.hr)
    if args.hr:
        kw["dam_column"] = (200.0, 14.0)
    part = cases.floodplain(px=px, py=py, rank=rank, **kw)
    serial = cases.floodplain(**kw)
    part.dt = serial.dt  # the same step on every rank (the serial case's)
    if args.recon:
        # second order: a fully wet, moving state (set by global cell id so that every rank sees the same field)
        part.config.reconstruction = serial.config.reconstruction = args.recon
        for c in (part, serial):
            g = c.mesh.array("cell_global_ids")[c.mesh.array("owned_to_local")].astype(np.float64)
            c.u0[:, 1] = 0.3 * np.sin(0.37 * g) * c.u0[:, 0]
            c.u0[:, 2] = 0.3 * np.cos(0.23 * g) * c.u0[:, 0]
    if args.mixed_paths:
        for c in (part, serial):
            m = c.mesh
            o2l = m.array("owned_to_local")
            g = m.array("cell_global_ids")[o2l].astype(np.float64)
            xc = m.array("cell_xc")[o2l]
            c.sources[:, 1] = np.where(xc < 0.25 * args.nx * 10.0 / 1.0, 1e-4 * np.sin(0.37 * g), 0.0)  # inside rank 0's block (px >= 2)
    op = Operator.from_case(part, device=local, halo=part.mesh.halo())
    if args.kernel:
        op.set_kernel(args.kernel, args.tile)
    if args.max_ctas:
        op.set_max_ctas(args.max_ctas)
    uid = [Operator.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    op.init_comm(world, rank, uid[0])
    if args.p2p:
        blobs = [None] * world
        dist.all_gather_object(blobs, op.halo_export(rank))
        op.halo_import(world, rank, blobs)
    m0 = torch.tensor([op.mass()], dtype=torch.float64, device="cuda")
    # three calls: each one primes the neighbours again, and the epochs of the fused push carry over between calls
    n1 = (args.steps - 1) // 2
    step = (lambda n: op.step_rk4(2 * serial.dt, n)) if args.rk4 else (lambda n: op.step_euler(serial.dt, n))
    step(n1)
    step(args.steps - 1 - n1)
    op.reset_diagnostics()  # the oracle's Courant diagnostic covers its last RHS evaluation only
    step(1)
    op.synchronize()
    u = op.get_state()
    m1 = torch.tensor([op.mass()], dtype=torch.float64, device="cuda")
    cmax = torch.tensor([op.courant().max_courant_num], dtype=torch.float64, device="cuda")
    dist.all_reduce(m0); dist.all_reduce(m1); dist.all_reduce(cmax, op=dist.ReduceOp.MAX)

    # gather (global id, state) on rank 0
    n_own = part.mesh.num_owned_cells
    gids = part.mesh.array("cell_global_ids")[part.mesh.array("owned_to_local")].astype(np.int64)
    sizes = [None] * world
    dist.all_gather_object(sizes, n_own)
    nmax = max(sizes)
    buf = torch.zeros((nmax, 4), dtype=torch.float64, device="cuda")
    buf[:n_own, 0] = torch.from_numpy(gids.astype(np.float64)).cuda()
    buf[:n_own, 1:] = torch.from_numpy(u).cuda()
    allbuf = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(allbuf, buf)
    ok = True
    if rank == 0:
        from util import assert_state_close, make_oracle

        oracle = make_oracle(serial)
        if args.rk4:
            oracle.step_rk4(2 * serial.dt, args.steps)
        else:
            oracle.step_euler(serial.dt, args.steps)
        sg = serial.mesh.array("cell_global_ids")[serial.mesh.array("owned_to_local")]
        u_ref_by_gid = np.zeros((int(sg.max()) + 1, 3))
        u_ref_by_gid[sg] = oracle.array("u_global").reshape(-1, 3)
        u_by_gid = np.full_like(u_ref_by_gid, np.nan)
        for r in range(world):
            b = allbuf[r][: sizes[r]].cpu().numpy()
            u_by_gid[b[:, 0].astype(np.int64)] = b[:, 1:]
        try:
            assert not np.isnan(u_by_gid[sg]).any(), "some global cells are owned by no rank"
            rel, _, mism = assert_state_close(u_by_gid[sg], u_ref_by_gid[sg], what=f"{world}-rank floodplain")
            assert abs(cmax.item() - oracle.courant()) <= 1e-12 * oracle.courant(), "max Courant number differs"
            print(f"MULTIGPU_PARITY ok ranks={world} p2p={args.p2p} hr={args.hr} kernel={args.kernel} steps={args.steps} max_rel={rel:.2e} wetdry_mismatch={mism} "
                  f"mass_drift={(m1.item() - m0.item()) / m0.item():.2e} courant={cmax.item():.4f} max_ctas={args.max_ctas} [{op.kernel_info().split(';')[0]}]", flush=True)
            if args.mixed_paths:
                assert "K4" not in op.kernel_info(), "rank 0 was expected to leave the K4 path (momentum sources)"
        except AssertionError as exc:
            ok = False
            print(f"MULTIGPU_PARITY FAILED: {exc}", flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    op.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1 else 1)


if __name__ == "__main__":
    main()
