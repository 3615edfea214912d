#!/usr/bin/env python
"""BASELINE.json configs[4]: a batch of synthetic 1024x1024 RGB images, every colour plane coded on its own (brute force,
8x8 ranges) and decoded again, image-parallel over the GPUs of one box: each rank takes a contiguous share of whole images,
encodes its planes one after the other (a plane fills the GPU), decodes all of them with ONE batched call (one CTA per
plane) and the winner records are all-gathered with NCCL.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29541 \
      profiles/run_batch.py [images] [side]
"""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fic_b200  # noqa: E402
from conftest import synthetic_image  # noqa: E402
from fic_b200.parallel import image_shares  # noqa: E402

n_images = int(sys.argv[1]) if len(sys.argv) > 1 else 256
side = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
mine = image_shares(n_images, rank, world)
share = max(len(image_shares(n_images, r, world)) for r in range(world))
# three distinct planes, shifted per image: the content is synthetic anyway, the work per plane is what counts
base = np.stack([synthetic_image(100 + k, side, side) for k in range(3)])
planes = np.stack([np.roll(base[c], 8 * i, axis=0) for i in mine for c in range(3)]) if mine else np.zeros((0, side, side), np.uint8)
h = fic_b200.Handle(local)
n = (side // 8) ** 2
pad_planes = share * 3


def run():
    idx, so, R = fic_b200.bf_encode_batch(planes, mode=1, handle=h)
    dec = fic_b200.bf_decode_batch(idx, so, side, side, 32, handle=h)
    if world > 1:
        t_idx = torch.zeros((pad_planes, n, 4), dtype=torch.int32, device=dev)
        t_so = torch.zeros((pad_planes, n, 2), dtype=torch.float64, device=dev)
        t_idx[:len(idx)] = torch.from_numpy(idx).to(dev)
        t_so[:len(so)] = torch.from_numpy(so).to(dev)
        g_idx = [torch.zeros_like(t_idx) for _ in range(world)]
        g_so = [torch.zeros_like(t_so) for _ in range(world)]
        dist.all_gather(g_idx, t_idx)
        dist.all_gather(g_so, t_so)
        torch.cuda.synchronize()
    return idx, dec


fic_b200.bf_encode(planes[0], mode=1, handle=h)      # warm-up: allocations
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
idx, dec = run()
dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
mse = float(np.mean((dec[0].astype(np.float64) - planes[0].astype(np.float64)) ** 2))
if rank == 0:
    sec = float(dt.item())
    print(json.dumps({"config": 4, "what": f"{n_images} synthetic {side}x{side} RGB images, per-plane brute-force encode + 32-sweep decode, "
                                           f"image-parallel over {world} GPU(s)", "n_gpus": world, "seconds": sec,
                      "images_per_s": n_images / sec, "planes_per_rank": len(planes),
                      "encode_matches_per_s": n_images * 3 * n * (side - 16) ** 2 * 8 / sec,
                      "psnr_plane0_db": float(10 * np.log10(255.0 ** 2 / mse))}), flush=True)
if world > 1:
    dist.destroy_process_group()
