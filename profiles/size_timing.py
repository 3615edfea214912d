#!/usr/bin/env python
"""Device-resident brute-force encode time of a synthetic square image: python profiles/size_timing.py 1024 [reps]"""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fic_b200  # noqa: E402
from conftest import synthetic_image  # noqa: E402

n = int(sys.argv[1])
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
img = synthetic_image(5, n, n)
h = fic_b200.Handle(0)
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(device=dev)
torch.cuda.set_stream(st)
nr = (n // 8) ** 2
d_img = torch.from_numpy(img).to(dev)
d_idx = torch.zeros((nr, 4), dtype=torch.int32, device=dev)
d_so = torch.zeros((nr, 2), dtype=torch.float64, device=dev)
d_R = torch.zeros((nr,), dtype=torch.float64, device=dev)
def run():
    fic_b200.bf_encode_dev(h, d_img.data_ptr(), n, n, d_idx.data_ptr(), d_so.data_ptr(), d_R.data_ptr(), flags=0, stream=st.cuda_stream)
run(); run()
torch.cuda.synchronize()
h.reset_counters()
t = time.perf_counter()
for _ in range(reps):
    run()
torch.cuda.synchronize()
dt = (time.perf_counter() - t) / reps
c = h.counters()
print(f"{n}x{n}: {dt * 1e3:.3f} ms/encode, {nr * (n - 16) ** 2 * 8 / dt:.4g} matches/s, survivors/range {c['survivors'] / reps / nr:.1f}, rescanned {c['rescan_ranges'] // reps}")
