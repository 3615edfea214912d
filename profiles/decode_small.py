"""Brute-force decode of small images (shared-memory shadow path): time and equality with the oracle.
Run on the GPU box: python profiles/decode_small.py"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fic_b200
from conftest import synthetic_image
from oracle import oracle as orc
imgs = np.load(os.path.join(ROOT, "tests/golden/images.npz"))
for name, img in (("lena256", imgs["lena256"]), ("synthetic 320x320", synthetic_image(3, 320, 320)),
                  ("synthetic 128x128", synthetic_image(4, 128, 128))):
    H, W = img.shape
    idx, so, R = fic_b200.bf_encode(img, mode=1)
    best = 1e9
    for _ in range(5):
        t = time.perf_counter(); d = fic_b200.bf_decode(idx, so, W, H, 32); best = min(best, time.perf_counter() - t)
    few = 1e9
    for _ in range(5):
        t = time.perf_counter(); fic_b200.bf_decode(idx, so, W, H, 2); few = min(few, time.perf_counter() - t)
    print(name, "2 sweeps: %.3f ms ->" % (few * 1e3), "%.1f us per further sweep;" % ((best - few) / 30 * 1e6), end=" ")
    print(name, "decode 32 sweeps: %.3f ms" % (best * 1e3), "equals oracle:", bool(np.array_equal(d, orc.bf_decode(idx, so, W, H, 32))))
