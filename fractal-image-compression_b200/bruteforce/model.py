"""Result containers of the brute-force codec.

Field names and order follow the reference (bruteforce/model.py:4-19): a transform is
``(start_i, start_j, orientation, rotation, s, o)`` and ``EncodedImage.domains`` lists one per
range block in row-major range order (bruteforce/bf_encoder.py:21-25).  The flat-array
converters are what the C ABI (include/fic.h: fic_bf_encode / fic_bf_decode) exchanges.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np


@dataclass
class Domain:
    start_i: int
    start_j: int
    orientation: int
    rotation: int
    s: float
    o: float


class EncodedImage:
    def __init__(self, width: int, height: int, range_size: int) -> None:
        self.width = width
        self.height = height
        self.range_size = range_size
        self.domains: list[Domain] = []

    # ---- flat <-> object form -------------------------------------------------------------
    @classmethod
    def from_arrays(cls, width: int, height: int, range_size: int, idx: np.ndarray, so: np.ndarray) -> "EncodedImage":
        enc = cls(width, height, range_size)
        enc.domains = [Domain(int(a), int(b), int(c), int(d), float(s), float(o))
                       for (a, b, c, d), (s, o) in zip(idx.tolist(), so.tolist())]
        return enc

    def to_arrays(self) -> tuple[np.ndarray, np.ndarray]:
        n = len(self.domains)
        idx = np.empty((n, 4), np.int32)
        so = np.empty((n, 2), np.float64)
        for k, d in enumerate(self.domains):
            idx[k] = (d.start_i, d.start_j, d.orientation, d.rotation)
            so[k] = (d.s, d.o)
        return idx, so

    def __len__(self) -> int:
        return len(self.domains)
