"""RGBQuadtreeEncoder 1024x1024x3, class API, median of 7: python profiles/rgb_qt_timing.py"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import synthetic_image  # noqa: E402
from quadtree.encoder import QuadtreeEncoder  # noqa: E402
from quadtree.rgb import RGBQuadtreeEncoder  # noqa: E402
rgb = np.ascontiguousarray(np.stack([synthetic_image(100 + k, 1024, 1024) for k in range(3)]).transpose(1, 2, 0))
qp = (7, 5, 1.0, False, False, 4, 8, 2.0)
out = {}
for mode in ("4:4:4", "4:2:0"):
    enc = RGBQuadtreeEncoder(QuadtreeEncoder(*qp), QuadtreeEncoder(*qp))
    enc.encode(rgb, mode)
    t = []
    for _ in range(7):
        t0 = time.perf_counter(); enc.encode(rgb, mode); t.append((time.perf_counter() - t0) * 1e3)
    out[mode] = round(float(np.median(t)), 3)
print(out)
