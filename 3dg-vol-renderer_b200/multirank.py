"""Host-side logic of the one-process-per-GPU mode, shared by bench.py and the tests.

Forward rendering has NO data-path collective: every sample depends only on (x, y, si), the scene is
replicated, and rank r renders the image rows y with (y - y0) % world == r
(vr_render_params.part_rank / part_count).  The only exchange is the optional assembly of the full
image on every rank (gather_image).

Inverse rendering has ONE real exchange per iteration: the per-Gaussian loss differences
F_diff[S, N] (inverse_integrator.h:157-179) are sums over pixels and views, so they decompose over
disjoint pixel sets; each rank accumulates its share in int64 fixed point (2^-32 units, device
buffer vr_sfd_fdiff_dev) and a SUM all-reduce of S*N int64 makes the total -- exact and independent
of rank order -- available everywhere; every rank then applies the same deterministic Adam step.

Works with any torch.distributed backend: "nccl" on the GPU box, "gloo" in the CPU tests.
"""
from __future__ import annotations

import numpy as np

FIXED_ONE = float(1 << 32)   # scale of the fixed-point F_diff sums (vr_inverse.cu kFixedScale)


def owned_rows(rank: int, world: int, y0: int, y1: int) -> range:
    """Image rows rank `rank` of `world` renders inside the window [y0, y1)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad row partition")
    return range(y0 + rank, y1, world)


def views_of_rank(rank: int, world: int, n_views: int) -> list:
    """Inverse mode with several target views: views are dealt round-robin; when there are fewer views
    than ranks every rank takes every view and the rows are split instead (see split_for_views)."""
    return list(range(rank, n_views, world)) if n_views >= world else list(range(n_views))


def split_for_views(rank: int, world: int, n_views: int):
    """-> (views, part_rank, part_count) for this rank: whole views when there are enough of them,
    otherwise all views with interleaved rows."""
    if n_views >= world:
        return views_of_rank(rank, world, n_views), 0, 1
    return list(range(n_views)), rank, world


def gather_image(local_img, rank: int, world: int, y0: int = 0, y1: int | None = None, group=None):
    """Every rank passes its [H, W, 3] float32 image in which only its owned rows are valid; returns the
    full image (valid on every rank).  Rows are disjoint, so this is a copy, not a sum: the result is
    bit-identical to a single-rank render."""
    import torch
    import torch.distributed as dist
    t = torch.as_tensor(local_img)
    if world == 1:
        return t
    H = t.shape[0]
    y1 = H if y1 is None else y1
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t.contiguous(), group=group)
    out = t.clone()
    for r in range(world):
        rows = torch.as_tensor(list(owned_rows(r, world, y0, y1)), dtype=torch.long, device=t.device)
        if rows.numel():
            out[rows] = parts[r][rows]
    return out


def allreduce_fdiff_(fixed, group=None):
    """In-place SUM all-reduce of the int64 fixed-point F_diff buffer (torch tensor, any device)."""
    import torch
    import torch.distributed as dist
    if fixed.dtype != torch.int64:
        raise TypeError("F_diff partial sums must be int64 fixed point")
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(fixed, op=dist.ReduceOp.SUM, group=group)
    return fixed


def fixed_to_float(fixed) -> np.ndarray:
    return np.asarray(fixed, dtype=np.int64).astype(np.float64) / FIXED_ONE


def float_to_fixed(x) -> np.ndarray:
    """Round-to-nearest conversion used by the device (__double2ll_rn)."""
    return np.rint(np.asarray(x, dtype=np.float64) * FIXED_ONE).astype(np.int64)


def sfd_gradient(F, signs, eps) -> np.ndarray:
    """grad[i] = mean_s F[s, i // 11] * signs[s, i] / eps[i]   (inverse_integrator.h:182-198), float64 -> float32."""
    F = np.asarray(F, np.float64); signs = np.asarray(signs, np.float64); eps = np.asarray(eps, np.float64)
    S, D = signs.shape
    g = (np.repeat(F, 11, axis=1)[:, :D] * signs / eps[None, :]).sum(axis=0) / S
    return g.astype(np.float32)


def device_int64_view(ptr: int, count: int, device_index: int):
    """Wrap a raw device pointer (vr_sfd_fdiff_dev) as a torch int64 CUDA tensor without copying, so that
    NCCL can all-reduce the context's own buffer."""
    import torch

    class _Holder:
        pass
    h = _Holder()
    h.__cuda_array_interface__ = {"shape": (count,), "typestr": "<i8", "data": (int(ptr), False), "version": 3, "strides": None}
    return torch.as_tensor(h, device=torch.device("cuda", device_index))
