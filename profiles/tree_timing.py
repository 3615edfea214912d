"""Wall time of the tree encoders through the C ABI (host image in, host node stream out), lena 512x512:
   python profiles/tree_timing.py   -> one JSON line (median of 7 calls after a warm-up)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
import fic_b200  # noqa: E402

img = np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))["lena512"]
h = fic_b200.default_handle(0)


def med(fn, reps=7):
    fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(t))


out = {
    "hv_512_ms": med(lambda: fic_b200.hv_encode(img, 4, 4, 16, 16, 5, 7, 25.0, handle=h), 5),
    "qt_512_ms": med(lambda: fic_b200.qt_encode(img, 7, 5, 1.0, False, False, 4, 8, 2.0, handle=h)),
    "qt_512_full_ms": med(lambda: fic_b200.qt_encode(img, 7, 5, 1.0, True, False, 4, 8, 2.0, handle=h)),
    "qt_512_dense_full_ms": med(lambda: fic_b200.qt_encode(img, 7, 5, 1.0, True, False, 4, 8, 2.0, 1, 1, handle=h), 3),
    "qt_512_dense_ms": med(lambda: fic_b200.qt_encode(img, 7, 5, 1.0, False, False, 4, 8, 2.0, 1, 1, handle=h)),
}
print(json.dumps(out))
