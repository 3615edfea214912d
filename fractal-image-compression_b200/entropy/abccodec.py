"""Encoder interface of the entropy stage (reference: entropy/abccodec.py)."""
from __future__ import annotations

from abc import ABC, abstractmethod

Code = int
CodeWidth = int


class EntropyEncoder(ABC):
    @abstractmethod
    def encode(self, val: int) -> tuple[Code, CodeWidth]:
        """(code word, its width in bits) for one symbol."""


class IdentityEncoder(EntropyEncoder):
    """Fixed-width pass-through used for dimensions that are not worth entropy coding."""

    def __init__(self, bit_width: CodeWidth) -> None:
        self.bit_width = bit_width

    def encode(self, val: int) -> tuple[Code, CodeWidth]:
        return val, self.bit_width
