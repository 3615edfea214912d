"""`BruteForceEncoder`: NumPy-float64 semantics of the reference (bruteforce/bf_encoder.py:7-71),
computed by the exact CUDA scan of libfic_b200.so (mode FIC_BF_MODE_NUMPY: no clamp of `s`,
threshold 1e-4, uint8-wrapped sum of range squares)."""
from __future__ import annotations

import numpy as np

import fic_b200
from bruteforce.model import EncodedImage


class BruteForceEncoder:
    def __init__(self, img_raw: np.ndarray, range_size=8, domain_size=16, strides=(1, 1)) -> None:
        """img_raw: 2-d uint8 array whose sides are multiples of `range_size` (reference docstring:
        "2d square array with side divisible by 16")."""
        self.img = img_raw
        self.range_size = range_size
        self.domain_size = domain_size
        self.strides = strides

    def encode(self) -> EncodedImage:
        height, width = self.img.shape
        idx, so, _ = fic_b200.bf_encode(self.img, self.range_size, self.domain_size, self.strides,
                                        mode=fic_b200.BF_MODE_NUMPY)
        if len(idx) and (height - self.domain_size <= 0 or width - self.domain_size <= 0):
            # bf_encoder.py:24 unpacks `None` when the lattice is empty
            raise TypeError("cannot unpack non-iterable NoneType object")
        return EncodedImage.from_arrays(width, height, self.range_size, idx, so)
