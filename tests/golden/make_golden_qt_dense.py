"""Golden fixture for the quadtree stress row (dense pools, full class search) at sizes the oracle needs minutes for.

    python tests/golden/make_golden_qt_dense.py            # writes tests/golden/qt_dense_full_golden.json

Runs the CPU oracle (oracle/, pinned to the reference by tests/test_oracle_golden.py) on
  * synthetic 272x272, seed 61, (7, 5, 1.0, True, True, 4, 8, 6.0, 1, 1)      -- about a minute
  * lena 512x512,               (7, 5, 1.0, True, False, 4, 8, 2.0, 1, 1)     -- the row of DESIGN section 6, tens of minutes
and stores the SHA-256 of the node table, the float64 values and the number of roots, plus the leaf count.
The GPU test (tests/test_tree_gpu.py::test_quadtree_dense_full_class_golden) hashes the CUDA result the same way."""
import hashlib
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from conftest import synthetic_image  # noqa: E402
from oracle import oracle  # noqa: E402


def digest(result):
    nodes, vals, n_roots = result
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(nodes, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(vals, dtype=np.float64).tobytes())
    h.update(int(n_roots).to_bytes(4, "little"))
    return h.hexdigest()


def main():
    oracle.lib()
    out_path = os.path.join(HERE, "qt_dense_full_golden.json")
    out = json.load(open(out_path)) if os.path.exists(out_path) else {}
    cases = {
        "synthetic272_seed61_positive": (synthetic_image(61, 272, 272), (7, 5, 1.0, True, True, 4, 8, 6.0, 1, 1)),
        "lena512_stress_row": (np.load(os.path.join(HERE, "images.npz"))["lena512"], (7, 5, 1.0, True, False, 4, 8, 2.0, 1, 1)),
    }
    for name, (img, params) in cases.items():
        if name in out and "--force" not in sys.argv:
            continue
        t = time.time()
        res = oracle.qt_encode(img, *params)
        out[name] = {"params": list(params), "shape": list(img.shape), "sha256": digest(res), "nodes": int(len(res[0])),
                     "oracle_seconds": round(time.time() - t, 1)}
        json.dump(out, open(out_path, "w"), indent=1, sort_keys=True)
        print(name, out[name], flush=True)


if __name__ == "__main__":
    main()
