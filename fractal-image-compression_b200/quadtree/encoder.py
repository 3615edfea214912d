"""`QuadtreeEncoder` (reference: quadtree/encoder.py:34-384) on the B200.

The constructor signature and the returned `QuadtreeImage` are the reference's.  The work --
integral-image classification of every domain and range (`_classify`), the 3x24 class pools
(`_build_domain_list`), the per-level batched search with the float32-sequential scorer
(`_find_min_err_domain`, `_find_optimal_mean_square_error`, `_calc_mean_square_error`) and the
split decisions -- happens in `fic_qt_encode` (csrc/tree_encode.cu).
"""
from __future__ import annotations

import numpy as np

import fic_b200
from quadtree.common import EncodingInfo, QuadtreeImage

NUM_MEAN_CLASSES = 3
NUM_VARIANCE_CLASSES = 24
POSITIVE = 1
NEGATIVE = -1
ZERO_TOLERANCE = 1e-6
SAME_AS_SIDE = -1


class QuadtreeEncoder:
    def __init__(self, scale_bits: int, offset_bits: int, max_scale: float, full_class_search: bool,
                 only_positive_scale: bool, min_range_side: int, max_range_side: int, error_tolerance: float,
                 domain_hstep: int = SAME_AS_SIDE, domain_vstep: int = SAME_AS_SIDE) -> None:
        self.scale_bits = scale_bits
        self.offset_bits = offset_bits
        self.max_scale = max_scale
        self.full_class_search = full_class_search
        self.only_positive_scale = only_positive_scale
        self.min_range_side = min_range_side
        self.max_range_side = max_range_side
        self.min_range_exponent = int(np.log2(min_range_side))
        self.max_range_exponent = int(np.log2(max_range_side))
        self.error_tolerance = error_tolerance
        self.domain_hstep = domain_hstep
        self.domain_vstep = domain_vstep
        # filled by encode(), kept for parity with the reference's attributes (encoder.py:65-75)
        self.max_square_exponent = None
        self.min_needed_range_exponent = None
        self.last_stream = None
        self._domain_lists = None
        self._pool_source = None      # what `domain_lists` is built from when somebody asks for it

    def encode(self, img: np.ndarray) -> QuadtreeImage:
        return self._encode(img, 0, img.shape)

    def encode_device(self, img_ptr: int, height: int, width: int) -> QuadtreeImage:
        """`encode` of a uint8 [height, width] plane that already lives in device memory (the colour wrapper's planes)."""
        return self._encode(None, img_ptr, (height, width))

    @property
    def domain_lists(self):
        """The reference's attribute of the same name (encoder.py:71-75, 261-283): per domain size -- exponents
        `min_needed_range_exponent + 1 .. max_range_exponent + 1` -- the 3 x 24 class lists of `(start_i, start_j, rotation,
        orientation)` rows, `[(0, 0, 0, 0)]` for an empty class.  The search itself uses the pools the library builds in
        device memory; this view is made on first access from the same classification (`fic_qt_classify`)."""
        if self._domain_lists is None and self._pool_source is not None:
            img = self._pool_source
            lists = []
            for e in range(self.min_needed_range_exponent + 1, self.max_range_exponent + 2):
                side = 2 ** e
                hstep = side if self.domain_hstep == SAME_AS_SIDE else self.domain_hstep
                vstep = side if self.domain_vstep == SAME_AS_SIDE else self.domain_vstep
                nv, nh = 1 + (img.shape[0] - side) // vstep, 1 + (img.shape[1] - side) // hstep
                ii, jj = np.meshgrid(np.arange(max(nv, 0)) * vstep, np.arange(max(nh, 0)) * hstep, indexing="ij")
                blocks = np.stack([ii.ravel(), jj.ravel(), np.full(ii.size, side), np.zeros(ii.size, np.int64)], 1)
                cls = fic_b200.qt_classify(img, blocks.astype(np.int32)) if len(blocks) else np.zeros((0, 4), np.int32)
                per_class = [[[] for _ in range(NUM_VARIANCE_CLASSES)] for _ in range(NUM_MEAN_CLASSES)]
                for (i, j, _, _), (mc, vc, rot, orient) in zip(blocks.tolist(), cls.tolist()):
                    per_class[mc][vc].append((i, j, rot, orient))
                lists.append([[np.array(c if c else [(0, 0, 0, 0)], np.int32) for c in row] for row in per_class])
            self._domain_lists = lists
        return self._domain_lists

    def _encode(self, img, img_ptr: int, shape) -> QuadtreeImage:
        height, width = shape
        self.max_square_exponent = min(int(np.log2(height)), int(np.log2(width)))
        self.min_needed_range_exponent = min(self.min_range_exponent,
                                             max(int(np.log2(_smallest_range_size(width, height))), 1))
        self._domain_lists = None
        self._pool_source = None if img is None else np.ascontiguousarray(img)
        nodes, vals, n_roots = fic_b200.qt_encode(
            img, self.scale_bits, self.offset_bits, self.max_scale, self.full_class_search, self.only_positive_scale,
            2 ** self.min_range_exponent, 2 ** self.max_range_exponent, self.error_tolerance, self.domain_hstep,
            self.domain_vstep, img_ptr=img_ptr, shape=shape)
        self.last_stream = (nodes, vals, n_roots)
        info = EncodingInfo(self.scale_bits, self.offset_bits, self.max_scale, width, height)
        return QuadtreeImage(info, stream=(nodes, vals, n_roots))   # node objects are built on first access


def _smallest_range_size(width: int, height: int) -> int:
    """encoder.py:81-95: the smallest square the ragged partition of a width x height image contains."""
    if width == 0:
        return height
    if height == 0:
        return width
    if width == 1 or height == 1:
        return 1
    max_square = 2 ** int(min(np.log2(width), np.log2(height)))
    if max_square == width and max_square == height:
        return max_square
    fits_on_x, fits_on_y = width // max_square, height // max_square
    return min(_smallest_range_size(fits_on_x * max_square, height - fits_on_y * max_square),
               _smallest_range_size(width - fits_on_x * max_square, height))


def _classify(img: np.ndarray, start_i: int, start_j: int, side: int) -> tuple[int, int, int, int]:
    """Mean class, variance class, rotation, orientation of one block (reference `_classify`, encoder.py:195-259).
    A signed int32 image (the reference passes ``-(img.astype(int32))`` for negative scales) is classified
    through the `negate` flag of `fic_qt_classify`."""
    a = np.asarray(img)
    neg = 0
    if a.dtype != np.uint8:
        if (a <= 0).all():
            a, neg = (-a).astype(np.uint8), 1
        else:
            a = a.astype(np.uint8)
    mc, vc, rot, orient = fic_b200.qt_classify(a, [(start_i, start_j, side, neg)])[0].tolist()
    return mc, vc, rot, orient
