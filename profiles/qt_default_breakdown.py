"""Default quadtree encode of lena 512x512: wall time of the C-ABI call against the Python wrapper's, and the kernels' share
(run under `ncu --metrics gpu__time_duration.sum` for the latter).  python profiles/qt_default_breakdown.py"""
import ctypes as C
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
import fic_b200  # noqa: E402

img = np.ascontiguousarray(np.load(os.path.join(ROOT, "tests", "golden", "images.npz"))["lena512"])
h = fic_b200.default_handle(0)
lib = fic_b200.lib()
cap = 2 * 512 * 512 + 64
nodes = np.empty((cap, 8), np.int32)
vals = np.empty((cap, 3), np.float64)
nodes[:] = 0; vals[:] = 0          # touch the pages once
nroots = C.c_int32(0)


def call():
    return lib.fic_qt_encode(h.ptr, img.ctypes.data_as(C.c_void_p), 512, 512, 7, 5, C.c_double(1.0), 0, 0, 4, 8, C.c_double(2.0), -1, -1,
                             nodes.ctypes.data_as(C.c_void_p), vals.ctypes.data_as(C.c_void_p), cap, C.byref(nroots), 0, None)


def med(fn, reps=30):
    for _ in range(3):
        fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter(); fn(); t.append((time.perf_counter() - t0) * 1e3)
    return round(float(np.median(t)), 4), round(min(t), 4)


print("C ABI, preallocated outputs (median, min ms):", med(call))
print("python wrapper qt_encode            :", med(lambda: fic_b200.qt_encode(img, 7, 5, 1.0, False, False, 4, 8, 2.0, handle=h)))
