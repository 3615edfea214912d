import os, sys, time
import numpy as np
ROOT = os.getcwd()
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT):
    sys.path.insert(0, p)
import fic_b200
img = np.load(os.path.join(ROOT, "tests/golden/images.npz"))["lena512"]
h = fic_b200.default_handle()
for mode in (0, 1):
    fic_b200.bf_encode(img, mode=mode)
    h.reset_counters()
    t = time.perf_counter(); a = fic_b200.bf_encode(img, mode=mode); t1 = time.perf_counter() - t
    c = h.counters()
    t = time.perf_counter(); b = fic_b200.bf_encode(img, mode=mode, flags=fic_b200.BF_EXACT_ONLY); t2 = time.perf_counter() - t
    print("mode", mode, "screen %.2f ms" % (t1 * 1e3), "exact %.1f ms" % (t2 * 1e3), "equal", all(np.array_equal(x, y) for x, y in zip(a, b)), c["survivors"], c["rescan_ranges"])
