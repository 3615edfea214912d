"""`JitedBruteForceEncoder`: the reference's numba semantics (bruteforce/jit_bfencoder.py:8-95) --
float32 sequential sums, `s` clamped to [-1, 1] -- on the B200.  For 8x8 ranges at stride 1 the
search runs as pool kernel -> tcgen05 screen GEMM -> exact rescoring; other shapes use the exact scan.
Both are bit-exact to the reference (tests/test_bf_gpu.py)."""
from __future__ import annotations

import numpy as np

import fic_b200
from bruteforce.model import EncodedImage


class JitedBruteForceEncoder:
    def __init__(self, img_raw: np.ndarray, range_size: int = 8, domain_size: int = 16) -> None:
        self.img = img_raw
        self.range_size = range_size
        self.domain_size = domain_size

    def encode(self) -> tuple[EncodedImage, list[np.float64]]:
        height, width = self.img.shape
        idx, so, R = fic_b200.bf_encode(self.img, self.range_size, self.domain_size, (1, 1),
                                        mode=fic_b200.BF_MODE_JIT)
        encoded = EncodedImage.from_arrays(width, height, self.range_size, idx, so)
        with np.errstate(invalid="ignore"):
            errs = list(np.sqrt(R))  # jit_bfencoder.py:28
        return encoded, errs
