"""Coding dimensions and the entropy-processor interface (reference: entropy/processor.py)."""
from __future__ import annotations

from enum import Enum


class CodingDimension(Enum):
    # values are list indices: 0..n-1 (processor.py:11-16)
    DOMAIN_ROW = 0
    DOMAIN_COL = 1
    SCALE = 2
    OFFSET = 3
