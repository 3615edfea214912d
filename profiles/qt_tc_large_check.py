"""One-off cross-checks of the quadtree tensor-core search against the oracle at sizes between the unit tests and the
512x512 golden: full class search with the default pools (384x448) and with steps (4, 4) (256x320)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "fractal-image-compression_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
from conftest import synthetic_image  # noqa: E402
from oracle import oracle  # noqa: E402
import fic_b200  # noqa: E402
oracle.lib()
ok = True
for shape, params, seed in (((384, 448), (7, 5, 1.0, True, False, 4, 8, 2.0), 91), ((256, 320), (7, 5, 1.0, True, False, 4, 8, 1.5, 4, 4), 92),
                            ((320, 256), (6, 6, 0.75, True, True, 4, 8, 2.5, 2, 3), 93)):
    img = synthetic_image(seed, *shape)
    img[10:40, 30:90] = 99
    t0 = time.time(); want = oracle.qt_encode(img, *params); t1 = time.time()
    got = fic_b200.qt_encode(img, *params)
    same = got[2] == want[2] and np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    ok = ok and same
    print(shape, params, "oracle %.1f s" % (t1 - t0), len(want[0]), "nodes", "equal" if same else "DIFFERENT")
print("OK" if ok else "FAILED")
