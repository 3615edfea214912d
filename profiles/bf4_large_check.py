"""4x4 brute force on a large image: the tensor-core path against the library's exact scan on slices of the ranges
(1024x768 synthetic; the exact scan needs ~50 us per range and position row, so only slices are compared)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import synthetic_image  # noqa: E402
import fic_b200  # noqa: E402
img = synthetic_image(77, 768, 1024)
img[100:140, 200:260] = 31
h = fic_b200.default_handle(0)
t0 = time.perf_counter(); full = fic_b200.bf_encode(img, range_size=4, domain_size=8, handle=h); t1 = time.perf_counter()
print("all", len(full[0]), "ranges in", round((t1 - t0) * 1e3, 1), "ms")
ok = True
for mode in (1, 0):
    for b, e in ((0, 64), (6400, 6464), (25 * 256 + 50, 25 * 256 + 70), (49152 - 64, 49152)):
        a = fic_b200.bf_encode(img, range_size=4, domain_size=8, mode=mode, range_begin=b, range_end=e, handle=h)
        x = fic_b200.bf_encode(img, range_size=4, domain_size=8, mode=mode, range_begin=b, range_end=e, flags=fic_b200.BF_EXACT_ONLY, handle=h)
        same = all(np.array_equal(p, q) for p, q in zip(a, x))
        if mode == 1:
            same = same and all(np.array_equal(p[b:e] if p.ndim else p, q) for p, q in zip(full, a))
        ok = ok and same
        print("mode", mode, "ranges", b, e, "equal" if same else "DIFFERENT")
print("OK" if ok else "FAILED")
