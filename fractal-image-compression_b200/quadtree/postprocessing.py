"""Post-decode filters of the quadtree codec (reference: quadtree/postprocessing.py:9-71).

`DeblockingQuadtreePostprocessor` blends the first row / column of every leaf block -- and the line just past it --
with the line beyond, using one weight pair for blocks of the deepest level and another for larger ones.  The
reference walks the forest leaf by leaf and lets later leaves overwrite earlier ones; `fic_qt_deblock`
(csrc/decode.cu) reproduces that order from the block geometry, one thread per pixel, and is bit-identical to it.
"""
from __future__ import annotations

from abc import ABC, abstractmethod

import numpy as np

import fic_b200
from quadtree.common import QuadtreeImage, forest_to_leaves


class QuadtreePostprocessor(ABC):
    @abstractmethod
    def postprocess(self, decoded_img: np.ndarray, quadtree_img: QuadtreeImage) -> np.ndarray:
        pass


class DeblockingQuadtreePostprocessor(QuadtreePostprocessor):
    def __init__(self) -> None:
        self.own_weight_max_depth = 5. / 6.
        self.neigh_weight_max_depth = 1. - self.own_weight_max_depth
        self.own_weight_not_max_depth = 2. / 3.
        self.neigh_weight_not_max_depth = 1. - self.own_weight_not_max_depth
        self.max_depth_range_size = 4
        self.res = None

    def _weights(self):
        return (self.own_weight_max_depth, self.neigh_weight_max_depth, self.own_weight_not_max_depth,
                self.neigh_weight_not_max_depth)

    def has_default_settings(self) -> bool:
        """True when the filter can run fused into the decoder (FIC_QT_DEBLOCK uses the reference's constants)."""
        return type(self) is DeblockingQuadtreePostprocessor and \
            self._weights() == (5. / 6., 1. - 5. / 6., 2. / 3., 1. - 2. / 3.) and self.max_depth_range_size == 4

    def postprocess(self, decoded_img: np.ndarray, quadtree_img: QuadtreeImage) -> np.ndarray:
        leaves, _ = forest_to_leaves(quadtree_img)
        self.res = fic_b200.qt_deblock(np.asarray(decoded_img, np.float64), leaves, self._weights(),
                                       int(self.max_depth_range_size))
        return self.res
