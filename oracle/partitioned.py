"""The reference's PARALLEL algorithm on the CPU oracle: one Operator per partition, ghosts filled from their owners before
every RHS evaluation.  TEST INFRASTRUCTURE ONLY (see swe_oracle.h): tests/, bench.py's CPU arm / cpu_baseline and bench.py's
parity check may import it, the product package never does.

What it restates: OperatorRHSFunction (src/rdysetup.c:1120-1139) on P ranks — DMGlobalToLocal (INSERT_VALUES) fills the one
overlap layer of src/rdydm.c:145-157, then ApplyOperator runs on u_local; interior edges that touch a ghost are evaluated
redundantly by both owners, there is no reverse scatter (src/swe/swe_petsc.c:272-274).  The exchange is a memcpy between
the parts of one process (`InProcessParts`, the "memcpy halo" CPU baseline of SURVEY section 8d) or a gather over
torch.distributed between processes (`DistributedPart`, the checker of the multi-GPU bench lines).
"""
from __future__ import annotations

import threading
import time

import numpy as np

from . import binding


def _fill(oracle, case):
    oracle.array("material_properties")[:] = case.mannings
    if case.sources is not None:
        oracle.array("external_sources")[:] = case.sources.ravel()
    for b, vals in case.dirichlet.items():
        oracle.array("boundary_values", b)[:] = np.asarray(vals).ravel()
    oracle.array("u_global")[:] = case.u0.ravel()


class Part:
    """One rank: its case (mesh block + halo plan) and its oracle."""

    def __init__(self, case, kind=None):
        self.case = case
        self.kind = kind or binding.best_kind()
        arr, nb = case.boundary_views()
        self.view = case.mesh.view()
        self.oracle = binding.Oracle(self.kind, case.config, self.view, arr if nb else None)
        self._keep = (arr,)
        _fill(self.oracle, case)
        h = case.mesh.halo_arrays()
        va = case.mesh.view_arrays()
        l2o = va["cell_local_to_owned"]
        self.peer_ranks = [int(r) for r in h["peer_ranks"]]
        so, ro = h["send_offsets"], h["recv_offsets"]
        # owned index (row of u_global) of every cell this rank sends, local index (row of u_local) of every ghost it receives
        self.send_owned = [l2o[h["send_cells"][so[p]:so[p + 1]]].astype(np.int64) for p in range(len(self.peer_ranks))]
        self.recv_local = [h["recv_cells"][ro[p]:ro[p + 1]].astype(np.int64) for p in range(len(self.peer_ranks))]
        self.num_owned = case.mesh.num_owned_cells
        # ghosts whose owner is never instantiated (bounded samples) keep the initial state of those cells; every instantiated owner
        # overwrites its part before the first RHS.  The initial state of a ghost comes from the same generator formula as an owned cell.
        if case.info.get("u0_ghost") is not None:
            self.u_local()[self.num_owned:] = case.info["u0_ghost"]

    def u_global(self):
        return self.oracle.array("u_global").reshape(-1, 3)

    def u_local(self):
        return self.oracle.array("u_local").reshape(-1, 3)

    def send_values(self, p):
        return self.u_global()[self.send_owned[p]]

    def set_ghosts(self, p, values):
        self.u_local()[self.recv_local[p]] = values

    def step(self, dt):
        self.oracle.rhs(dt)
        self.oracle.euler_update(dt)


class InProcessParts:
    """All P parts of one mesh in this process, one host thread per part (ctypes releases the GIL)."""

    def __init__(self, parts):
        self.parts = parts
        self.by_rank = {p.case.info["rank"]: p for p in parts}
        self.cells_per_step = sum(p.num_owned for p in parts)

    def _pull_ghosts(self, part):
        me = part.case.info["rank"]
        for p, r in enumerate(part.peer_ranks):
            other = self.by_rank.get(r)
            if other is None:  # the owner lies outside this bounded sample of the mesh: its ghosts stay frozen at their initial value
                continue
            q = other.peer_ranks.index(me)
            part.set_ghosts(p, other.send_values(q))

    def step_euler(self, dt, nsteps):
        """nsteps Euler steps on every part; returns wall seconds (ghost exchange included)."""
        n = len(self.parts)
        bar = threading.Barrier(n)
        errors = []

        def work(part):
            try:
                for _ in range(nsteps):
                    bar.wait()  # every part finished its update: u_global is final
                    self._pull_ghosts(part)
                    bar.wait()  # every part has read its neighbours: they may update now
                    part.step(dt)
            except Exception as exc:  # pragma: no cover - a broken barrier must not hang the others
                errors.append(exc)
                bar.abort()

        threads = [threading.Thread(target=work, args=(p,)) for p in self.parts]
        t0 = time.perf_counter()
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return time.perf_counter() - t0

    def state_by_global_id(self):
        gmax = max(int(p.case.mesh.array("cell_global_ids").max()) for p in self.parts)
        out = np.full((gmax + 1, 3), np.nan)
        for p in self.parts:
            m = p.case.mesh
            gids = m.array("cell_global_ids")[m.array("owned_to_local")]
            out[gids] = p.u_global()
        return out


class DistributedPart:
    """This process' part of a mesh partitioned over torch.distributed ranks (one per GPU in bench.py).  The ghost exchange gathers
    every rank's border values with all_gather_object: small (a few thousand cells per neighbour) and backend-agnostic."""

    def __init__(self, case, dist, rank, kind=None):
        self.part = Part(case, kind)
        self.dist = dist
        self.rank = rank

    def _exchange(self):
        part = self.part
        mine = {r: part.send_values(p).copy() for p, r in enumerate(part.peer_ranks)}
        everyone = [None] * self.dist.get_world_size()
        self.dist.all_gather_object(everyone, mine)
        for p, r in enumerate(part.peer_ranks):
            part.set_ghosts(p, everyone[r][self.rank])

    def step_euler(self, dt, nsteps):
        for _ in range(nsteps):
            self._exchange()
            self.part.step(dt)

    def u_global(self):
        return self.part.u_global()

    def courant(self):
        return self.part.oracle.courant()
