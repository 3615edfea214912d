#!/usr/bin/env python
"""Achieved HBM bandwidth of the byte kernels next to the quadtree codec (colour planes, deblocking filter), device buffers,
CUDA events on the launching stream, L2 flushed between repetitions.  Run on the GPU box:
    python profiles/post_kernels.py > gpurun_out/post_kernels.jsonl
Algorithmic bytes: rgb -> planes 3 B read + 3 B (4:4:4) or 1.5 B (4:2:0) written per pixel, and the reverse; deblocking
8 B read + 8 B written per pixel (float64 in and out; the 4-byte leaf index per pixel and the neighbours are extra traffic)."""
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

peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
PEAK = float(peaks.get("hbm_gbs", 0)) or None
dev = torch.device("cuda", 0)
h = fic_b200.default_handle(0)
lib = fic_b200.lib()
stream = torch.cuda.Stream(dev)          # a real stream: the library maps a NULL stream to its own
torch.cuda.set_stream(stream)
sp = C.c_void_p(stream.cuda_stream)
assert stream.cuda_stream != 0
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
vp = C.c_void_p


def timed(fn, reps=10):
    fn()
    stream.synchronize()
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
    gbs = nbytes / ms / 1e6
    print(json.dumps(dict(kernel=what, ms=ms, algorithmic_bytes=nbytes, achieved_gbs=gbs, peak_gbs=PEAK,
                          frac=(gbs / PEAK if PEAK else None), **kw)), flush=True)


N = M = 8192
npx = N * M
rgb = torch.randint(0, 256, (N, M, 3), dtype=torch.uint8, device=dev)
y = torch.empty((N, M), dtype=torch.uint8, device=dev)
for sub, name in ((0, "4:4:4"), (1, "4:2:0")):
    cs = (N // 2, M // 2) if sub else (N, M)
    cb, cr = torch.empty(cs, dtype=torch.uint8, device=dev), torch.empty(cs, dtype=torch.uint8, device=dev)
    out = torch.empty_like(rgb)

    def fwd():
        rc = lib.fic_rgb_to_ycbcr(h.ptr, vp(rgb.data_ptr()), N, M, sub, vp(y.data_ptr()), vp(cb.data_ptr()), vp(cr.data_ptr()),
                                  fic_b200.DEVICE_PTRS, sp)
        assert rc == 0

    def bwd():
        rc = lib.fic_ycbcr_to_rgb(h.ptr, vp(y.data_ptr()), vp(cb.data_ptr()), vp(cr.data_ptr()), N, M, sub, vp(out.data_ptr()),
                                  fic_b200.DEVICE_PTRS, sp)
        assert rc == 0

    planes = npx + 2 * cs[0] * cs[1]
    emit(f"rgb_to_ycbcr {name} {N}x{M}", timed(fwd), 3 * npx + planes)
    emit(f"ycbcr_to_rgb {name} {N}x{M}", timed(bwd), 3 * npx + planes)

# deblocking on a 4096 x 4096 float64 image tiled by 8x8 and 4x4 leaves
N = M = 4096
side = np.where((np.add.outer(np.arange(N // 8), np.arange(M // 8)) % 2) == 0, 8, 4)
leaves = []
for bi in range(N // 8):
    for bj in range(M // 8):
        if side[bi, bj] == 8:
            leaves.append((8 * bi, 8 * bj, 8, 8, 0, 0, 0, 0))
        else:
            leaves += [(8 * bi + a, 8 * bj + b, 4, 4, 0, 0, 0, 0) for a in (0, 4) for b in (0, 4)]
leaves = torch.from_numpy(np.array(leaves, np.int32)).to(dev)
img = torch.rand((N, M), dtype=torch.float64, device=dev) * 255
res = torch.empty_like(img)


def deblock():
    rc = lib.fic_qt_deblock(h.ptr, vp(img.data_ptr()), vp(leaves.data_ptr()), leaves.shape[0], N, M, None, -1,
                            vp(res.data_ptr()), fic_b200.DEVICE_PTRS, sp)
    assert rc == 0


emit(f"fic_qt_deblock {N}x{M}, {leaves.shape[0]} leaves (paint + keys + filter)", timed(deblock), 16 * N * M,
     note="whole call: leaf-index map memset + paint, root list upload, key kernel, filter kernel")
