#!/usr/bin/env python
"""Host<->device copy ceiling of the node for the e2e figure of bench.py.

Every rank (one per GPU, torchrun) copies the bytes one `rdyb_operator_apply_host` call moves - 240 MB up and 240 MB down for
the 10 M-cell block - from / to pinned host memory with one cudaMemcpyAsync per direction on two streams, concurrently with all
other ranks, and reports the time of such a "step".  bench.py's e2e (which also runs the kernel and the AoS<->SoA passes between
the copies) can be read as a fraction of this ceiling: cells / (ceiling ms).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29655 tools/pcie_ceiling.py
"""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cells = 9999392
    n = 3 * cells
    h_in = torch.zeros(n, dtype=torch.float64).pin_memory()
    h_out = torch.zeros(n, dtype=torch.float64).pin_memory()
    d_in = torch.empty(n, dtype=torch.float64, device="cuda")
    d_out = torch.zeros(n, dtype=torch.float64, device="cuda")
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
    out = {}
    for mode in ("duplex", "h2d_only", "d2h_only"):
        def step():
            if mode != "d2h_only":
                with torch.cuda.stream(s_up):
                    d_in.copy_(h_in, non_blocking=True)
            if mode != "h2d_only":
                with torch.cuda.stream(s_dn):
                    h_out.copy_(d_out, non_blocking=True)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        k = 10
        t0 = time.perf_counter()
        for _ in range(k):
            step()
        torch.cuda.synchronize()
        t = (time.perf_counter() - t0) / k
        if world > 1:
            tt = torch.tensor([t], dtype=torch.float64, device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            t = float(tt.item())
        out[mode] = {"ms_per_step": 1e3 * t, "GBps_per_direction_per_gpu": 8e-9 * n / t,
                     "cell_updates_per_s_ceiling": world * cells / t if mode == "duplex" else None}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "bytes_per_direction_per_gpu": 8 * n, "pinned": True, "copies": "one cudaMemcpyAsync per direction, two streams", **out}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
