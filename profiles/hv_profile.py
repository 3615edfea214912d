#!/usr/bin/env python
"""One HV encode of lena512 (BASELINE.json configs[2]); run under `ncu --metrics gpu__time_duration.sum` for the launch list."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT):
    sys.path.insert(0, p)
import fic_b200  # noqa: E402

img = np.load(os.path.join(ROOT, "tests/golden/images.npz"))["lena512"]
hp = (4, 4, 16, 16, 5, 7, 25.0)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 1):
    t = time.perf_counter()
    out = fic_b200.hv_encode(img, *hp)
    print("hv_encode", time.perf_counter() - t, "s", len(out[0]), "nodes", file=sys.stderr)
