"""`HVEncoder` (reference: hvv2/encoder.py:10-188) on the B200: adaptive rectangular ranges, domain lattice of
step (h, w), four shape-preserving isometries.  Partitioning is level-synchronous on the device
(`fic_hv_encode`, csrc/tree_encode.cu): one launch searches every pending rectangle of a tree depth, another
computes the biased-difference split of those that must be divided."""
from __future__ import annotations

import numpy as np

import fic_b200
from hvv2.common import HVImage
from quadtree.common import EncodingInfo


class HVEncoder:
    def __init__(self, min_range_height: int, min_range_width: int, max_range_height: int, max_range_width: int,
                 scale_bits: int, offset_bits: int, err_threshold: float, max_scale: float = 1.0) -> None:
        self.min_range_height = min_range_height
        self.min_range_width = min_range_width
        self.max_range_height = max_range_height
        self.max_range_width = max_range_width
        self.err_threshold = err_threshold
        self.scale_bits = scale_bits
        self.offset_bits = offset_bits
        self.max_scale = max_scale
        self.last_stream = None

    def encode(self, img: np.ndarray) -> HVImage:
        nodes, vals = fic_b200.hv_encode(img, self.min_range_height, self.min_range_width, self.max_range_height,
                                         self.max_range_width, self.scale_bits, self.offset_bits, self.err_threshold,
                                         self.max_scale)
        self.last_stream = (nodes, vals)
        info = EncodingInfo(scale_bits=self.scale_bits, offset_bits=self.offset_bits, max_scale=self.max_scale,
                            img_width=img.shape[1], img_height=img.shape[0])
        return HVImage(info=info, stream=(nodes, vals))   # node objects are built on first access
