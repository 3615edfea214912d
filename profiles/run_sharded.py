#!/usr/bin/env python
"""BASELINE.json configs[3]: brute-force encode of ONE synthetic 4096x4096 image with its range blocks sharded over
the GPUs of one box (contiguous range slices, no data-path collective; the winner records are all-gathered with NCCL).

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29531 \
      profiles/run_sharded.py [side]
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
from fic_b200.parallel import range_slice  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
img = synthetic_image(0, side, side)
n = (side // 8) ** 2
b, e = range_slice(n, rank, world)
cnt = e - b
pad = max(range_slice(n, r, world)[1] - range_slice(n, r, world)[0] for r in range(world))
h = fic_b200.Handle(local)
st = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(st)
d_img = torch.from_numpy(img).to(dev)
d_idx = torch.zeros((pad, 4), dtype=torch.int32, device=dev)
d_so = torch.zeros((pad, 2), dtype=torch.float64, device=dev)
d_R = torch.zeros((pad,), dtype=torch.float64, device=dev)
g_idx = [torch.zeros_like(d_idx) for _ in range(world)]
g_so = [torch.zeros_like(d_so) for _ in range(world)]


def run():
    fic_b200.bf_encode_dev(h, d_img.data_ptr(), side, side, d_idx.data_ptr(), d_so.data_ptr(), d_R.data_ptr(),
                           range_begin=b, range_end=e, stream=st.cuda_stream)
    if world > 1:
        dist.all_gather(g_idx, d_idx)
        dist.all_gather(g_so, d_so)


run()                      # warm-up (allocations, NCCL set-up)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
t0 = time.perf_counter()
run()
torch.cuda.synchronize()
dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
if world > 1:
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
if rank == 0:
    sec = float(dt.item())
    c = h.counters()
    print(json.dumps({"config": 3, "what": f"bruteforce-jit synthetic {side}x{side}, {n} ranges sharded over {world} GPU(s)",
                      "n_gpus": world, "seconds": sec, "matches_per_s": n * (side - 16) ** 2 * 8 / sec,
                      "ranges_per_rank": cnt, "rank0_rescan_ranges": c["rescan_ranges"]}), flush=True)
if world > 1:
    dist.destroy_process_group()
