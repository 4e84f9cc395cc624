"""N>1 host logic on CPU: two gloo ranks (world_size 2) exercise the row partition, the image
assembly and the exact fixed-point all-reduce of the inverse path (3dg-vol-renderer_b200/multirank.py).
The device work is replaced by a deterministic per-pixel function -- what is tested is the plumbing
the GPU ranks use unchanged (bench.py launches the same code over NCCL)."""
from __future__ import annotations

import os
import socket

import numpy as np
import pytest

from conftest import load_pkg_module


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _fake_render(W, H, rows):
    """stand-in for vr_render with a row partition: only the owned rows are written"""
    img = np.zeros((H, W, 3), np.float32)
    y = np.array(list(rows), np.int64)
    if y.size:
        x = np.arange(W)
        img[y] = (np.sin(0.37 * x[None, :, None] + 1.3 * y[:, None, None] + np.arange(3)[None, None, :]) ** 2).astype(np.float32)
    return img


def _worker(rank, world, port, out_dir):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mr = load_pkg_module("multirank")
    W, H = 37, 29
    # ---- forward: interleaved rows, assembled image == single-rank image, bit for bit
    local = _fake_render(W, H, mr.owned_rows(rank, world, 0, H))
    full = mr.gather_image(local, rank, world).numpy()
    assert np.array_equal(full, _fake_render(W, H, range(H)))
    # ---- inverse: per-rank partial F_diff in fixed point, SUM all-reduce, same gradient everywhere
    S, N = 3, 41
    rs = np.random.RandomState(7)
    per_pixel = rs.normal(size=(H, S, N)) * (rs.uniform(size=(H, S, N)) < 0.2)    # contribution of each image row
    mine = mr.float_to_fixed(per_pixel[list(mr.owned_rows(rank, world, 0, H))]).sum(axis=0)
    buf = torch.from_numpy(mine.reshape(-1).copy())
    mr.allreduce_fdiff_(buf)
    total = mr.float_to_fixed(per_pixel).sum(axis=0).reshape(-1)
    assert np.array_equal(buf.numpy(), total)          # exact: integer sums are order independent
    signs = np.where(rs.uniform(size=(S, 11 * N)) < 0.5, -1, 1).astype(np.int8)
    eps = np.tile(np.array([.02] * 3 + [.1] * 3 + [.05] * 3 + [.25, .5], np.float32), N)
    g = mr.sfd_gradient(mr.fixed_to_float(buf.numpy()).reshape(S, N), signs, eps)
    gs = [torch.empty(g.size, dtype=torch.float32) for _ in range(world)]
    dist.all_gather(gs, torch.from_numpy(g))
    assert all(torch.equal(gs[0], t) for t in gs)      # replicated deterministic step needs identical gradients
    # ---- view dealing
    assert mr.split_for_views(rank, world, 16) == (list(range(rank, 16, world)), 0, 1)
    assert mr.split_for_views(rank, world, 1) == ([0], rank, world)
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")


def test_two_gloo_ranks(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_row_partition_covers_every_row_once():
    mr = load_pkg_module("multirank")
    for world in (1, 2, 3, 8):
        for y0, y1 in ((0, 17), (5, 6), (3, 64)):
            rows = sorted(y for r in range(world) for y in mr.owned_rows(r, world, y0, y1))
            assert rows == list(range(y0, y1))
    with pytest.raises(ValueError):
        mr.owned_rows(2, 2, 0, 4)


def test_reference_arm_runs_on_rank0_only(tmp_path):
    """bench.py --impl reference under torchrun: ranks != 0 exit 0 without work or output."""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0 and r.stdout.strip() == ""
