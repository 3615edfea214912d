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
    """`width`, `height`, `range_size`, `domains` as in the reference (bruteforce/model.py:14-19).  An image that comes
    from the encoder keeps the flat record arrays; the `Domain` objects of `domains` are made when the list is first
    read (and the arrays are dropped then: the list may be edited), so encode -> decode does not pay for one Python
    object per range block."""

    def __init__(self, width: int, height: int, range_size: int) -> None:
        self.width = width
        self.height = height
        self.range_size = range_size
        self._domains: list[Domain] | None = []
        self._arrays: tuple[np.ndarray, np.ndarray] | None = None

    @property
    def domains(self) -> list[Domain]:
        if self._domains is None:
            idx, so = self._arrays
            self._domains = [Domain(a, b, c, d, s, o) for (a, b, c, d), (s, o) in zip(idx.tolist(), so.tolist())]
            self._arrays = None
        return self._domains

    @domains.setter
    def domains(self, value: list[Domain]) -> None:
        self._domains = value
        self._arrays = None

    # ---- flat <-> object form -------------------------------------------------------------
    @classmethod
    def from_arrays(cls, width: int, height: int, range_size: int, idx: np.ndarray, so: np.ndarray) -> "EncodedImage":
        enc = cls(width, height, range_size)
        enc._domains = None
        enc._arrays = (np.ascontiguousarray(idx, np.int32).reshape(-1, 4), np.ascontiguousarray(so, np.float64).reshape(-1, 2))
        return enc

    def to_arrays(self) -> tuple[np.ndarray, np.ndarray]:
        if self._arrays is not None:
            return self._arrays
        doms = self._domains
        idx = np.array([(d.start_i, d.start_j, d.orientation, d.rotation) for d in doms], np.int32).reshape(-1, 4)
        so = np.array([(d.s, d.o) for d in doms], np.float64).reshape(-1, 2)
        return idx, so

    def __len__(self) -> int:
        return self._arrays[0].shape[0] if self._arrays is not None else len(self._domains)
