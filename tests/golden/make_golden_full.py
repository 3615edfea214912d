#!/usr/bin/env python
"""Full-size brute-force fixtures made by the CPU oracle (oracle/fic_oracle.c, itself pinned bit for bit to the
unmodified reference by tests/test_oracle_golden.py):

  * the SHA-256 of ALL 4 096 records (idx int32[n,4] | so float64[n,2] | R float64[n], C order, concatenated) of the
    JitedBruteForceEncoder encode of lena 512x512 -- the benchmark's image;
  * 16 records of the synthetic 4096x4096 image of BASELINE.json configs[3] (conftest.synthetic_image(0, 4096, 4096)),
    ranges spread over the image.

    python tests/golden/make_golden_full.py          # ~10 min on 8 host cores
Writes tests/golden/bf_full_golden.npz.
"""
import hashlib
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import synthetic_image  # noqa: E402
from oracle import oracle as orc  # noqa: E402


def records_sha(idx, so, R) -> str:
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(idx, np.int32).tobytes())
    h.update(np.ascontiguousarray(so, np.float64).tobytes())
    h.update(np.ascontiguousarray(R, np.float64).tobytes())
    return h.hexdigest()


def main():
    out = {}
    big = synthetic_image(0, 4096, 4096)
    n = (4096 // 8) ** 2
    sel = np.sort(np.random.default_rng(4096).choice(n, 16, replace=False)).astype(np.int64)
    t0 = time.time()
    def one(r):   # the oracle's OpenMP loop runs over ranges, so scattered ranges go to a thread pool, one thread each
        idx, so, R = orc.bf_encode(big, mode=1, range_begin=int(r), range_end=int(r) + 1, nthreads=1)
        print("4096 range", int(r), idx[0], so[0], R, f"{time.time() - t0:.0f}s", flush=True)
        return np.concatenate([idx[0].astype(np.float64), so[0], R])

    from concurrent.futures import ThreadPoolExecutor
    with ThreadPoolExecutor(os.cpu_count() or 4) as ex:
        recs = list(ex.map(one, sel))
    out["syn4096_sel"] = sel
    out["syn4096_rec"] = np.stack(recs)
    out["syn4096_image_sha256"] = np.array(hashlib.sha256(big.tobytes()).hexdigest())

    lena = np.load(os.path.join(HERE, "images.npz"))["lena512"]
    t0 = time.time()
    idx, so, R = orc.bf_encode(lena, mode=1)
    print(f"lena512 full encode by the oracle: {time.time() - t0:.0f}s", flush=True)
    out["jit512_sha256"] = np.array(records_sha(idx, so, R))
    out["jit512_idx"] = idx          # 64 KB: also kept in the clear so a mismatch can be located
    np.savez_compressed(os.path.join(HERE, "bf_full_golden.npz"), **out)
    print("wrote bf_full_golden.npz", out["jit512_sha256"])


if __name__ == "__main__":
    main()
