"""Multi-GPU host logic: one process per GPU (torch.distributed), no data-path collective.

Range blocks are independent given the image, so a single large image is sharded by contiguous
range-block slices and a batch by whole images; the only exchange is an all-gather of the winner records
(16 B index + 16 B (s,o) + 8 B R per range).  The functions take the encode callable as a parameter so the same
slicing / gathering code is exercised on CPU (gloo, world size 2) in tests/test_parallel_cpu.py.
"""
from __future__ import annotations

from typing import Callable, Sequence

import numpy as np


def range_slice(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous row-major slice [begin, end) of `total` range blocks owned by `rank`: sizes differ by at most one."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def image_shares(n_images: int, rank: int, world: int) -> list[int]:
    """Indices of the batch images `rank` encodes (contiguous shares, like range_slice)."""
    b, e = range_slice(n_images, rank, world)
    return list(range(b, e))


def _gather_rows(local: np.ndarray, counts: Sequence[int], dist, device) -> np.ndarray:
    """all_gather of row blocks with per-rank row counts (padded to the largest block)."""
    import torch
    world = len(counts)
    width = int(np.prod(local.shape[1:])) if local.ndim > 1 else 1
    pad = max(counts)
    buf = torch.zeros((pad, width), dtype=torch.from_numpy(local.reshape(len(local), -1)[:0]).dtype, device=device)
    if len(local):
        buf[:len(local)] = torch.from_numpy(np.ascontiguousarray(local).reshape(len(local), -1)).to(device)
    out = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(out, buf)
    parts = [o[:c].cpu().numpy() for o, c in zip(out, counts)]
    return np.concatenate(parts).reshape((-1,) + local.shape[1:])


def encode_sharded(img: np.ndarray, encode_slice: Callable[[np.ndarray, int, int], tuple], range_size: int = 8,
                   device=None):
    """Brute-force encode of ONE image with its range blocks sharded over the ranks of the default process group.
    `encode_slice(img, begin, end)` -> (idx[n,4], so[n,2], R[n]) for that slice (fic_b200.bf_encode with range_begin /
    range_end on a GPU rank).  Every rank returns the full records in range order."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    H, W = img.shape
    total = (H // range_size) * (W // range_size)
    begin, end = range_slice(total, rank, world)
    idx, so, R = encode_slice(img, begin, end)
    counts = [range_slice(total, r, world)[1] - range_slice(total, r, world)[0] for r in range(world)]
    dev = device if device is not None else "cpu"
    return (_gather_rows(np.asarray(idx, np.int32), counts, dist, dev),
            _gather_rows(np.asarray(so, np.float64), counts, dist, dev),
            _gather_rows(np.asarray(R, np.float64), counts, dist, dev))


def encode_batch(imgs: np.ndarray, encode_image: Callable[[np.ndarray], tuple], device=None):
    """Image-parallel encode of a batch [B,H,W]: each rank encodes its share of whole images, records are gathered.
    Returns (idx[B,n,4], so[B,n,2], R[B,n]) on every rank."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    B = imgs.shape[0]
    mine = image_shares(B, rank, world)
    res = [encode_image(imgs[b]) for b in mine]
    n = None
    for r in res:
        n = len(r[0])
    if n is None:  # a rank without images still has to know the record count
        probe = encode_image(imgs[0])
        n = len(probe[0])
    idx = np.stack([r[0] for r in res]) if res else np.zeros((0, n, 4), np.int32)
    so = np.stack([r[1] for r in res]) if res else np.zeros((0, n, 2), np.float64)
    R = np.stack([r[2] for r in res]) if res else np.zeros((0, n), np.float64)
    counts = [len(image_shares(B, r, world)) for r in range(world)]
    dev = device if device is not None else "cpu"
    return (_gather_rows(idx.astype(np.int32), counts, dist, dev), _gather_rows(so, counts, dist, dev),
            _gather_rows(R, counts, dist, dev))


def decode_batch(idx: np.ndarray, so: np.ndarray, width: int, height: int,
                 decode_planes: Callable[[np.ndarray, np.ndarray], np.ndarray], device=None) -> np.ndarray:
    """Image-parallel decode of a batch of records idx[B,n,4], so[B,n,2]: each rank decodes its share of planes in one
    call (`decode_planes(idx_share, so_share)` -> uint8 [b,height,width]; fic_b200.bf_decode_batch on a GPU rank, which
    runs the planes concurrently), the pixels are gathered.  Returns uint8 [B,height,width] on every rank."""
    import torch.distributed as dist
    rank, world = dist.get_rank(), dist.get_world_size()
    B = idx.shape[0]
    mine = image_shares(B, rank, world)
    if mine:
        local = np.asarray(decode_planes(idx[mine[0]:mine[-1] + 1], so[mine[0]:mine[-1] + 1]), np.uint8)
    else:
        local = np.zeros((0, height, width), np.uint8)
    counts = [len(image_shares(B, r, world)) for r in range(world)]
    dev = device if device is not None else "cpu"
    return _gather_rows(local, counts, dist, dev)
