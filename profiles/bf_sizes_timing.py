"""Brute-force encode at range sizes 4, 8 and 16 (lena 256 / 512), wall time through the C ABI with host buffers:
   python profiles/bf_sizes_timing.py"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
import fic_b200  # noqa: E402

imgs = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))
h = fic_b200.default_handle(0)
out = {}
for name in ("lena256", "lena512"):
    img = imgs[name]
    for rs in (4, 8, 16):
        fn = lambda: fic_b200.bf_encode(img, range_size=rs, domain_size=2 * rs, handle=h)
        fn()
        t = []
        for _ in range(3):
            t0 = time.perf_counter(); fn(); t.append((time.perf_counter() - t0) * 1e3)
        H, W = img.shape
        matches = (H // rs) * (W // rs) * (H - 2 * rs) * (W - 2 * rs) * 8
        out[f"{name}_r{rs}"] = {"ms": round(min(t), 3), "matches_per_s": float(f"{matches / (min(t) * 1e-3):.4g}")}
print(json.dumps(out))
