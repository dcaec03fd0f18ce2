"""GPU: the N > 1 path (partitioned mesh, per-step NCCL halo exchange overlapped with interior cells) against the serial CPU
oracle.  Needs >= 2 GPUs; the worker is tests/multigpu_parity.py under torchrun (one process per GPU)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpus():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


def _run(nranks, extra, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "tests", "multigpu_parity.py"), *extra]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0 and "MULTIGPU_PARITY ok" in p.stdout, p.stdout[-3000:] + p.stderr[-3000:]
    return p.stdout


@pytest.mark.parametrize("kernel,tile", [(0, 0), (3, 224), (2, 256)])
def test_two_ranks_nccl_halo_matches_serial_oracle(kernel, tile):
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--kernel", str(kernel), "--tile", str(tile)], 29517 + kernel)


@pytest.mark.parametrize("kernel,tile", [(0, 0), (3, 224)])
def test_two_ranks_peer_memory_halo_matches_serial_oracle(kernel, tile):
    """The Euler fast path: border cells are written straight into the neighbour's ghost slots over NVLink (CUDA IPC)."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--p2p", "--kernel", str(kernel), "--tile", str(tile), "--steps", "80"], 29551 + kernel)


def test_two_ranks_hydrostatic_reconstruction():
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--hr", "--kernel", "3", "--tile", "224"], 29531)


def test_four_ranks_2x2_blocks():
    if _ngpus() < 4:
        pytest.skip("needs 4 GPUs")
    _run(4, ["--p2p", "--kernel", "3", "--tile", "224", "--nx", "128", "--ny", "96"], 29541)


@pytest.mark.parametrize("hr", [False, True], ids=["roe", "hr"])
def test_two_ranks_fused_push_many_tiles_per_cta(hr):
    """The fused in-kernel push with the grid capped at 4 CTAs: a CTA owns several frontier tiles AND many interior tiles, so the
    neighbour wait, the per-tile sends, the device counter and the epoch flag are exercised in the regime of the 10 M-cell bench."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    out = _run(2, ["--p2p", "--kernel", "3", "--tile", "224", "--nx", "192", "--ny", "128", "--steps", "90", "--max-ctas", "4"] + (["--hr"] if hr else []), 29581 + hr)
    assert "K4" in out


def test_two_ranks_on_different_step_paths_stay_in_step():
    """the peer-memory step path is a rank-local decision (tile shapes, momentum sources).  Rank 0 gets momentum
    sources and runs the general kernel + push kernel, rank 1 the fused-push K4; both advance the halo epoch by one priming push
    per call plus one per step, so three step_euler calls stay consistent."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--p2p", "--kernel", "3", "--tile", "224", "--steps", "80", "--mixed-paths"], 29587)


@pytest.mark.parametrize("kernel,tile", [(0, 0), (3, 224)])
def test_two_ranks_rk4_stages_over_the_peer_memory_halo(kernel, tile):
    """TSRK4 on a partitioned mesh without NCCL on the data path: each of the four RHS evaluations of a step pushes the border cells
    of its stage vector into the neighbour's ghost slots of the same (peer-mapped) buffer."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--p2p", "--rk4", "--kernel", str(kernel), "--tile", str(tile), "--steps", "40"], 29575 + kernel)


def test_four_ranks_fused_push_many_tiles_per_cta():
    if _ngpus() < 4:
        pytest.skip("needs 4 GPUs")
    _run(4, ["--p2p", "--kernel", "3", "--tile", "224", "--nx", "192", "--ny", "160", "--steps", "90", "--max-ctas", "5"], 29585)


@pytest.mark.parametrize("cap", [0, 6], ids=["fullgrid", "cap6"])
def test_eight_ranks_4x2_blocks(cap):
    """4 x 2 blocks: inner ranks have three neighbours (two faces and no corner: ghosts are face neighbours only).  The fused push's
    release / acquire protocol (stores, fence.sys, device counter, epoch flag per neighbour) with the most neighbours the bench uses."""
    if _ngpus() < 8:
        pytest.skip("needs 8 GPUs")
    _run(8, ["--p2p", "--kernel", "3", "--tile", "224", "--nx", "256", "--ny", "128", "--steps", "90"] + (["--max-ctas", str(cap)] if cap else []), 29590 + cap)


def test_two_ranks_second_order_exchanges_gradients():
    """MUSCL on a partitioned mesh: the ghost cells' gradients come from their owners (CommunicateCellGradients,
    src/operator_fluxes_ceed.c:1056-1098) over the same halo plan, after the state halo."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    _run(2, ["--recon", "1", "--kernel", "0", "--steps", "60"], 29561)


@pytest.mark.parametrize("tile", [128, 192])
def test_two_ranks_second_order_on_the_tiled_layout(tile):
    """The tile kernels on a partitioned mesh: frontier tiles list ghost cells in their halo and read the exchanged ghost gradients.
    128-cell tiles run K3M, 192-cell tiles K4M with the gradient array (a ghost in the first ring rules the in-kernel gradients out)."""
    if _ngpus() < 2:
        pytest.skip("needs 2 GPUs")
    out = _run(2, ["--recon", "1", "--kernel", "1", "--tile", str(tile), "--steps", "60"], 29571 + (tile == 192))
    assert ("K4M" in out and "muscl_gradient_kernel" in out) if tile == 192 else ("K3M" in out or "K2M" in out)
