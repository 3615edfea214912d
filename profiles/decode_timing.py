import os, sys, time
import numpy as np
ROOT = os.getcwd()
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import fic_b200
from conftest import synthetic_image
img = synthetic_image(7, 1024, 1024)
idx, so, R = fic_b200.bf_encode(img, mode=1)
for _ in range(2):
    t = time.perf_counter(); d = fic_b200.bf_decode(idx, so, 1024, 1024, 32); t1 = time.perf_counter() - t
print("single 1024^2 decode", t1 * 1e3, "ms")
I = np.stack([idx] * 3); S = np.stack([so] * 3)
for _ in range(2):
    t = time.perf_counter(); d3 = fic_b200.bf_decode_batch(I, S, 1024, 1024, 32); t3 = time.perf_counter() - t
print("3 planes", t3 * 1e3, "ms", np.array_equal(d3[0], d), np.array_equal(d3[2], d))
I = np.stack([idx] * 96); S = np.stack([so] * 96)
for _ in range(2):
    t = time.perf_counter(); d96 = fic_b200.bf_decode_batch(I, S, 1024, 1024, 32); t96 = time.perf_counter() - t
print("96 planes", t96 * 1e3, "ms", np.array_equal(d96[95], d))
