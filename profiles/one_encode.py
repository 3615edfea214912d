"""Two brute-force encodes of the bundled 512x512 image through the C ABI (the second one is what ncu captures):
   ncu --set full --import-source on --clock-control none -k regex:bf_screen_kernel --launch-skip 3 --launch-count 3 \\
       -o gpurun_out/screen python profiles/one_encode.py"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
import fic_b200  # noqa: E402

img = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))["lena512"]
h = fic_b200.default_handle(0)
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
    idx, so, R = fic_b200.bf_encode(img, handle=h)
print("ok", idx[:2].tolist(), h.counters())
