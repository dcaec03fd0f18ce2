"""CPU, world_size 2 over gloo: the host side of the N > 1 path — block partition, one layer of face-neighbour ghosts
(DMPlexDistributeOverlap(dm, 1), src/rdydm.c:145-157) and the halo plan that replaces the PetscSF of DMGlobalToLocal
(src/rdysetup.c:1133-1134).

Each process builds ITS block and exchanges states with torch.distributed send/recv following its RdybHaloView, exactly the
messages the CUDA path sends over NCCL (rdyb_operator_halo_exchange) — then
  * every ghost slot holds its owner's value (checked through global cell ids);
  * one oracle right-hand side per rank on the ghost-filled local vector equals the serial oracle's RHS on the owned cells
    (shared edges are evaluated redundantly by both ranks: no reverse scatter, src/swe/swe_petsc.c:272-274);
  * the RCB partitioner + extract_local give the same guarantees on a mesh that is not a block grid.
No GPU is involved; the oracle is the checker here, as everywhere in tests/.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _exchange(halo, rank, local_vals):
    """local_vals: [num_local, k] float64; fills the ghost rows from their owners following the halo plan"""
    reqs, bufs = [], []
    for p, peer in enumerate(halo["peer_ranks"]):
        s = halo["send_cells"][halo["send_offsets"][p]:halo["send_offsets"][p + 1]]
        r = halo["recv_cells"][halo["recv_offsets"][p]:halo["recv_offsets"][p + 1]]
        out = torch.from_numpy(np.ascontiguousarray(local_vals[s]))
        inc = torch.zeros((len(r), local_vals.shape[1]), dtype=torch.float64)
        bufs.append((r, inc))
        # lower rank sends first: gloo send/recv pairs must be matched in the same order on both sides
        if rank < int(peer):
            dist.send(out, int(peer)); dist.recv(inc, int(peer))
        else:
            dist.recv(inc, int(peer)); dist.send(out, int(peer))
    for r, inc in bufs:
        local_vals[r] = inc.numpy()
    return local_vals


def _worker(rank, world, port, mode, results):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from rdycore_b200 import cases
        from util import make_oracle

        serial = cases.floodplain(nx=24, ny=20, eta0=8.0)
        rng = np.random.default_rng(3)
        serial.u0[:, 1] = 0.2 * rng.normal(size=serial.u0.shape[0]) * serial.u0[:, 0]
        serial.u0[:, 2] = 0.2 * rng.normal(size=serial.u0.shape[0]) * serial.u0[:, 0]
        gid_serial = serial.mesh.array("cell_global_ids")
        ng = int(gid_serial.max()) + 1
        u_by_gid = np.zeros((ng, 3)); u_by_gid[gid_serial] = serial.u0
        if mode == "blocks":
            part = cases.floodplain(nx=24, ny=20, eta0=8.0, px=world, py=1, rank=rank)
            mesh = part.mesh
        else:  # RCB on the serial mesh, then extract this rank's local mesh
            p = serial.mesh.partition_rcb(world)
            counts = np.bincount(p, minlength=world)
            assert counts.max() - counts.min() <= 1, "RCB parts must be balanced"
            mesh = serial.mesh.extract_local(p, world, rank)
        halo = mesh.halo_arrays()
        gids = mesh.array("cell_global_ids")
        n_own, n_loc = mesh.num_owned_cells, mesh.num_cells
        va = mesh.view_arrays()
        assert va["cell_is_owned"][:n_own].all() and not va["cell_is_owned"][n_own:].any()  # owned first, ghosts after
        assert sorted(halo["recv_cells"].tolist()) == list(range(n_own, n_loc))              # every ghost is received exactly once
        assert (halo["send_cells"] < n_own).all()
        # 1. ghosts receive their owners' values
        local = np.full((n_loc, 4), np.nan)
        local[:n_own, :3] = u_by_gid[gids[:n_own]]
        local[:n_own, 3] = gids[:n_own]
        local = _exchange(halo, rank, local)
        assert not np.isnan(local).any()
        assert np.array_equal(local[:, 3].astype(np.int64), gids)
        assert np.array_equal(local[:, :3], u_by_gid[gids])
        # every cell of the mesh is owned by exactly one rank
        owned = [None] * world
        dist.all_gather_object(owned, gids[:n_own].tolist())
        allowned = np.sort(np.concatenate([np.asarray(o) for o in owned]))
        assert np.array_equal(allowned, np.sort(gid_serial))
        # 2. local oracle RHS on the ghost-filled vector == serial RHS on the owned cells
        if mode == "blocks":
            so = make_oracle(serial)
            F_by_gid = np.zeros((ng, 3)); F_by_gid[gid_serial] = so.rhs(serial.dt).reshape(-1, 3)
            po = make_oracle(part)
            po.array("u_global")[:] = local[:n_own, :3].ravel()
            ul = po.array("u_local").reshape(-1, 3)
            ul[n_own:] = local[n_own:, :3]           # the caller's job: ghosts (DMGlobalToLocal); rhs() fills the owned rows
            F = po.rhs(serial.dt).reshape(-1, 3)
            scale = np.abs(F_by_gid).max(axis=0)
            assert (np.abs(F - F_by_gid[gids[:n_own]]) / scale).max() < 1e-14
        results[rank] = "ok"
    except BaseException as e:  # noqa: BLE001
        results[rank] = f"{type(e).__name__}: {e}"
        raise
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("mode", ["blocks", "rcb"])
def test_two_ranks_over_gloo(mode):
    world = 2
    with mp.Manager() as m:
        results = m.dict()
        mp.spawn(_worker, args=(world, _free_port(), mode, results), nprocs=world, join=True)
        assert dict(results) == {0: "ok", 1: "ok"}
