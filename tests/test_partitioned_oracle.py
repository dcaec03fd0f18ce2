"""CPU: the partitioned CPU oracle (oracle/partitioned.py) — P reference operators on the blocks of one mesh with a ghost exchange
before every RHS (DMGlobalToLocal, src/rdysetup.c:1133-1134) — against the serial oracle.

It is the CPU arm of bench.py (`--impl reference`, `cpu_baseline`: the bench's own mesh cut into one block per host core, memcpy
halo) and the checker behind `checks.parity` of the multi-GPU bench lines (one process per GPU, ghosts through
torch.distributed), so it has to reproduce the serial result itself first: shared edges are evaluated redundantly from
bit-identical inputs (src/swe/swe_petsc.c:272-274), hence the partitioned state equals the serial one up to the summation order
inside a cell - which the local edge numbering of a block may change - i.e. to round-off, and the total mass is conserved.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle.partitioned import DistributedPart, InProcessParts, Part
from rdycore_b200 import cases

from util import assert_state_close, make_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
KW = dict(nx=40, ny=32, eta0=8.0)


def _serial_by_gid(steps, **kw):
    serial = cases.floodplain(**kw)
    o = make_oracle(serial)
    o.step_euler(serial.dt, steps)
    gids = serial.mesh.array("cell_global_ids")[serial.mesh.array("owned_to_local")]
    out = np.zeros((int(gids.max()) + 1, 3))
    out[gids] = o.array("u_global").reshape(-1, 3)
    return serial, out


@pytest.mark.parametrize("hr", [False, True])
@pytest.mark.parametrize("px,py", [(2, 1), (2, 2), (3, 2)])
def test_in_process_parts_reproduce_the_serial_oracle(px, py, hr):
    kw = dict(KW, hr=hr)
    if hr:
        kw.update(eta0=6.0, dam_column=(120.0, 12.0))
    serial, ref = _serial_by_gid(60, **kw)
    parts = []
    for r in range(px * py):
        c = cases.floodplain(px=px, py=py, rank=r, **kw)
        c.dt = serial.dt
        parts.append(Part(c))
    job = InProcessParts(parts)
    assert job.cells_per_step == serial.mesh.num_owned_cells
    job.step_euler(serial.dt, 60)
    u = job.state_by_global_id()
    assert not np.isnan(u).any()
    assert_state_close(u, ref, what=f"{px}x{py} in-process parts")
    assert np.abs(ref[:, 1]).max() > 1e-3  # the water moves: not a static comparison


def _worker(rank, world, port, results):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        serial, ref = _serial_by_gid(40, **KW)
        c = cases.floodplain(px=world, py=1, rank=rank, **KW)
        c.dt = serial.dt
        d = DistributedPart(c, dist, rank)
        d.step_euler(serial.dt, 40)
        gids = c.mesh.array("cell_global_ids")[c.mesh.array("owned_to_local")]
        assert_state_close(d.u_global(), ref[gids], what=f"rank {rank} of {world}")
        results[rank] = "ok"
    except BaseException as e:  # noqa: BLE001
        results[rank] = f"{type(e).__name__}: {e}"
        raise
    finally:
        dist.destroy_process_group()


def test_distributed_parts_over_gloo():
    world = 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    with mp.Manager() as m:
        results = m.dict()
        mp.spawn(_worker, args=(world, port, results), nprocs=world, join=True)
        assert dict(results) == {0: "ok", 1: "ok"}
