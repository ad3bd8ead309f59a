"""Multi-rank plumbing on the CPU (gloo, world_size 2): the view partition and the order-preserving
variable-length all-gather of the per-rank clouds (output must equal the 1-rank order)."""
import os
import socket
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import mod


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, counts, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        d = mod("distributed")
        g = torch.Generator().manual_seed(100 + rank)
        n = counts[rank]
        xyz = torch.rand(n, 3, generator=g)
        rgb = torch.rand(n, 3, generator=g)
        gx, gr, c = d.allgather_points(xyz, rgb)
        out[rank] = (gx.numpy().copy(), gr.numpy().copy(), c.numpy().copy())
    finally:
        dist.destroy_process_group()


def _gather_worker(rank, world, port, n, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        knn = mod("knn")
        perm = torch.randperm(n, generator=torch.Generator().manual_seed(5))   # "sorted position" -> original index
        a, b = knn.query_range(n, world, rank)
        mine = perm[a:b]
        rows = torch.stack([mine.float() * 2.0, mine.float() + 0.5], 1)       # a row is a function of its point
        full = knn.gather_rows_by_index(rows, mine.to(torch.int32), n)
        one = knn.gather_rows_by_index(rows[:, 0].contiguous(), mine, n)
        out[rank] = (full.numpy().copy(), one.numpy().copy(), (a, b))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [11, 8, 1])
def test_query_split_gather_restores_original_order_gloo_world2(n):
    """Host logic of the query-split post-processing (knn.local_outlier_factor_sharded): ranks own
    contiguous blocks of the sorted query positions, and the gathered per-point rows land at the
    points' original indices, ragged blocks and an empty rank included."""
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_gather_worker, args=(world, _free_port(), n, out), nprocs=world, join=True)
    want = torch.arange(n).float()
    ranges = [out[r][2] for r in range(world)]
    assert ranges[0][0] == 0 and ranges[-1][1] == n and ranges[0][1] == ranges[1][0]
    for r in range(world):
        full, one, _ = out[r]
        assert np.array_equal(full[:, 0], (want * 2).numpy()) and np.array_equal(full[:, 1], (want + 0.5).numpy())
        assert np.array_equal(one, (want * 2).numpy())


def test_view_partition_covers_views_in_order():
    d = mod("distributed")
    for n_views, world in [(200, 8), (7, 2), (3, 4), (0, 2), (300, 4)]:
        got = [i for r in range(world) for i in d.view_partition(n_views, world, r)]
        assert got == list(range(n_views))


def test_allgather_points_gloo_world2_is_rank_ordered_concat():
    world, counts = 2, [5, 9]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), counts, out), nprocs=world, join=True)
    exp_x, exp_r = [], []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)
        exp_x.append(torch.rand(counts[r], 3, generator=g)); exp_r.append(torch.rand(counts[r], 3, generator=g))
    exp_x, exp_r = torch.cat(exp_x).numpy(), torch.cat(exp_r).numpy()
    for r in range(world):
        gx, gr, c = out[r]
        assert np.array_equal(gx, exp_x) and np.array_equal(gr, exp_r) and c.tolist() == counts


def test_allgather_points_handles_an_empty_rank():
    world, counts = 2, [0, 4]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), counts, out), nprocs=world, join=True)
    for r in range(world):
        assert out[r][0].shape == (4, 3) and out[r][2].tolist() == counts


@pytest.mark.gpu
def test_fused_emit_exchange_matches_allgather_and_single_gpu():
    """2+ GPUs only (the round-end GPU box has one: skipped there; run by hand with gpurun --gpus 2)."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    n = min(torch.cuda.device_count(), 4)
    script = str(Path(__file__).resolve().parent / "exchange_check.py")
    for views in ("11", "1"):  # "1": every rank but the first has no view at all
        r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                            "--master-addr", "127.0.0.1", "--master-port", "29533", script],
                           capture_output=True, text=True, timeout=600, env={**os.environ, "MDI_CHECK_VIEWS": views})
        assert r.returncode == 0 and "exchange ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.gpu
def test_query_split_postprocess_is_bit_identical_to_single_gpu():
    """2+ GPUs only: LOF / extents with the queries split over ranks equal the single-GPU results."""
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    n = min(torch.cuda.device_count(), 4)
    script = str(Path(__file__).resolve().parent / "postprocess_check.py")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
                        "--master-addr", "127.0.0.1", "--master-port", "29534", script],
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "postprocess ok" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
