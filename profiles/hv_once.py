"""One HV encode of lena 512x512 (after a warm-up call) -- for `ncu --metrics gpu__time_duration.sum` launch lists."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
import fic_b200  # noqa: E402

img = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))["lena512"]
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
for _ in range(n):
    nodes, vals = fic_b200.hv_encode(img, 4, 4, 16, 16, 5, 7, 25.0)
print(len(nodes))
