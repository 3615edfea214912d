#!/usr/bin/env python
"""One quadtree encode of lena512 (BASELINE.json configs[1]) through the C ABI; run under ncu for the launch list."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT):
    sys.path.insert(0, p)
import fic_b200  # noqa: E402

img = np.load(os.path.join(ROOT, "tests/golden/images.npz"))["lena512"]
qp = (7, 5, 1.0, False, False, 4, 8, 2.0)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 1):
    t = time.perf_counter()
    out = fic_b200.qt_encode(img, *qp)
    print("qt_encode", (time.perf_counter() - t) * 1e3, "ms", len(out[0]), "nodes", file=sys.stderr)
