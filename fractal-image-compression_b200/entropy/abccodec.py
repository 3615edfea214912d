"""Encoder interface of the entropy stage (reference: entropy/abccodec.py).

Besides the reference's one-symbol `encode`, every encoder can translate a whole column of symbols at once
(`encode_column`): the file writers hand over the numpy column of a field and get the code words and their widths back
as arrays, ready for `BitBuffer.write_many`.
"""
from __future__ import annotations

from abc import ABC, abstractmethod

import numpy as np

Code = int
CodeWidth = int


class EntropyEncoder(ABC):
    @abstractmethod
    def encode(self, val: int) -> tuple[Code, CodeWidth]:
        """(code word, its width in bits) for one symbol."""

    def encode_column(self, values: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
        """(code words uint64, widths int64) of every symbol of `values`: one `encode` per DISTINCT symbol."""
        uniq, inverse = np.unique(np.asarray(values), return_inverse=True)
        pairs = [self.encode(int(v)) for v in uniq]
        codes = np.array([p[0] for p in pairs], np.uint64)
        widths = np.array([p[1] for p in pairs], np.int64)
        return codes[inverse], widths[inverse]


class IdentityEncoder(EntropyEncoder):
    """Fixed-width pass-through used for dimensions that are not worth entropy coding."""

    __slots__ = ("bit_width",)

    def __init__(self, bit_width: CodeWidth) -> None:
        self.bit_width = bit_width

    def encode(self, val: int) -> tuple[Code, CodeWidth]:
        return val, self.bit_width

    def encode_column(self, values: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
        v = np.asarray(values)
        return v.astype(np.uint64), np.full(v.shape, self.bit_width, np.int64)
