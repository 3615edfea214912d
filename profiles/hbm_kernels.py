#!/usr/bin/env python
"""The HBM-bound kernels either side of the correlation GEMM -- domain pooling and the decoders -- at sizes that leave
the 126 MB L2: one call of each, so that

    ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none --csv \\
        --log-file gpurun_out/r2_hbm_ncu.csv python profiles/hbm_kernels.py

lists every kernel with its DRAM bytes and duration (summarised by `profiles/ncu_summary.py hbm` into
profiles/r2_hbm_ncu.json), and, run without ncu, times the calls with CUDA events on the launching stream with the L2
flushed in between (profiles/r2_hbm_kernels.jsonl).

Algorithmic bytes: pooling  H*W (image) + NP*(128 + 8 + 8 + 32) written  (pool row, hk, two integer sums, eight
float sums per stride-1 position); tree decode  H*W*(4*8 read + 8 written + 4 leaf index) per sweep + 48 B per leaf;
final clip 8 B read + 1 B written per pixel; brute-force decode: the image lives in one CTA's shared memory or L2, so the
figure is wall time per sweep, not bandwidth.
"""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import fic_b200  # noqa: E402
from conftest import synthetic_image  # noqa: E402

peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
PEAK = float(peaks.get("hbm_gbs", 0)) or None
dev = torch.device("cuda", 0)
h = fic_b200.default_handle(0)
stream = torch.cuda.Stream(dev)
torch.cuda.set_stream(stream)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
quick = len(sys.argv) > 1 and sys.argv[1] == "once"      # under ncu: one call each, no repetitions


def timed(fn, reps=5):
    fn()
    stream.synchronize()
    if quick:
        return None
    ms = []
    for _ in range(reps):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        fn()
        b.record(stream)
        stream.synchronize()
        ms.append(a.elapsed_time(b))
    return float(np.median(ms))


def emit(what, ms, nbytes, **kw):
    if ms is None:
        return
    gbs = nbytes / ms / 1e6
    print(json.dumps(dict(call=what, ms=ms, algorithmic_bytes=nbytes, achieved_gbs=gbs, peak_gbs=PEAK,
                          frac=(gbs / PEAK if PEAK else None), **kw)), flush=True)


# ---- pooling: the first 32 ranges of a 4096 x 4096 image (pool + statistics over ALL 16.6 M positions, then a small screen)
side = 4096
img = torch.from_numpy(synthetic_image(0, side, side)).to(dev)
n = 32
d_idx = torch.zeros((n, 4), dtype=torch.int32, device=dev)
d_so = torch.zeros((n, 2), dtype=torch.float64, device=dev)
d_R = torch.zeros((n,), dtype=torch.float64, device=dev)
NP = (side - 16) ** 2
ms = timed(lambda: fic_b200.bf_encode_dev(h, img.data_ptr(), side, side, d_idx.data_ptr(), d_so.data_ptr(), d_R.data_ptr(),
                                          range_begin=0, range_end=n, stream=stream.cuda_stream))
emit("fic_bf_encode 4096x4096, 32 ranges (pool kernel + 32-range screen; see the ncu list for the pool kernel alone)", ms,
     side * side + NP * 176, positions=NP)
del img


# ---- tree decoders on synthetic leaf tables: every leaf an 8 x 8 (quadtree) / 8 x 4 (HV) block with a random domain
def leaf_table(H, W, lh, lw, rng):
    ii, jj = np.meshgrid(np.arange(0, H, lh), np.arange(0, W, lw), indexing="ij")
    k = ii.size
    L = np.zeros((k, 8), np.int32)
    L[:, 0], L[:, 1], L[:, 2], L[:, 3] = ii.ravel(), jj.ravel(), lh, lw
    L[:, 4] = rng.integers(0, H - 2 * lh + 1, k)
    L[:, 5] = rng.integers(0, W - 2 * lw + 1, k)
    L[:, 6] = rng.integers(0, 2, k)
    L[:, 7] = rng.integers(0, 4, k) if lh == lw else 2 * rng.integers(0, 2, k)
    so = np.stack([rng.uniform(-0.9, 0.9, k), rng.uniform(0, 255, k)], 1)
    return L, so


rng = np.random.default_rng(5)
for side in (4096, 8192):
    for kind, name, (lh, lw) in ((fic_b200.TREE_QUADTREE, "quadtree", (8, 8)), (fic_b200.TREE_HV, "hv", (8, 4))):
        L, so = leaf_table(side, side, lh, lw, rng)
        d_L = torch.from_numpy(L).to(dev)
        d_s = torch.from_numpy(so).to(dev)
        out = torch.empty((side, side), dtype=torch.uint8, device=dev)
        sweeps = 4
        ms = timed(lambda: fic_b200.tree_decode_dev(h, kind, d_L.data_ptr(), d_s.data_ptr(), len(L), side, side, out.data_ptr(),
                                                    iterations=sweeps, stream=stream.cuda_stream), reps=3)
        npx = side * side
        emit(f"fic_tree_decode {name} {side}x{side}, {len(L)} leaves, {sweeps} sweeps + clip", ms,
             sweeps * (npx * 44 + len(L) * 48) + npx * 9 + npx * 4, leaves=len(L))
        del d_L, d_s, out

# ---- deblocking filter, 4096 x 4096 float64, 8 x 8 leaves
side = 4096
L, _ = leaf_table(side, side, 8, 8, rng)
x = torch.rand((side, side), dtype=torch.float64, device=dev) * 255
res = torch.empty_like(x)
d_L = torch.from_numpy(L).to(dev)
lib = fic_b200.lib()
vp = C.c_void_p
ms = timed(lambda: fic_b200._check(lib.fic_qt_deblock(h.ptr, vp(x.data_ptr()), vp(d_L.data_ptr()), len(L), side, side, None, -1,
                                                      vp(res.data_ptr()), fic_b200.DEVICE_PTRS, vp(stream.cuda_stream)), "deblock"))
emit(f"fic_qt_deblock {side}x{side}, {len(L)} leaves", ms, side * side * 16, leaves=len(L))
del x, res, d_L

# ---- brute-force decoder, 1024 x 1024, 32 sweeps, random contractive records (one CTA per plane: latency, not bandwidth)
side = 1024
nr = (side // 8) ** 2
idx = np.zeros((3, nr, 4), np.int32)
idx[..., 0] = rng.integers(0, side - 16, (3, nr))
idx[..., 1] = rng.integers(0, side - 16, (3, nr))
idx[..., 2] = rng.integers(0, 2, (3, nr))
idx[..., 3] = rng.integers(0, 4, (3, nr))
so = np.stack([rng.uniform(-0.9, 0.9, (3, nr)), rng.uniform(0, 255, (3, nr))], -1)
import time
fic_b200.bf_decode_batch(idx, so, side, side, 32, handle=h)
if not quick:
    t0 = time.perf_counter()
    fic_b200.bf_decode_batch(idx, so, side, side, 32, handle=h)
    dt = (time.perf_counter() - t0) * 1e3
    print(json.dumps(dict(call="fic_bf_decode_batch 3 planes 1024x1024, 32 sweeps (host records in, host pixels out)", ms=dt,
                          ms_per_sweep=dt / 32)), flush=True)
