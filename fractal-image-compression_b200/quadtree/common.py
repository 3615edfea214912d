"""Data model of the quadtree codec (reference: quadtree/common.py:10-75).

Same class and field names as the reference so its serialisers, decoders and user code work on
these objects unchanged.  ``stream_to_forest`` / ``forest_to_leaves`` convert between the tree
objects and the flat arrays of the C ABI (include/fic.h: fic_qt_encode, fic_tree_decode).
"""
from __future__ import annotations

import gc
from dataclasses import dataclass, replace
from typing import Callable, Optional

import numpy as np

from entropy.processor import CodingDimension

NORMAL_ORIENTATION = 0
TRANSPOSED_ORIENTATION = 1
MAX_GRAY = 255


@dataclass(slots=True)
class Domain:
    start_i: int
    start_j: int
    orientation: int
    rotation: int
    scale: float          # pre-quantisation s (debug only in the reference)
    offset: float         # pre-quantisation o
    quantized_scale: int
    quantized_offset: int


@dataclass
class EncodingInfo:
    scale_bits: int
    offset_bits: int
    max_scale: float
    img_width: int
    img_height: int

    def copy(self) -> "EncodingInfo":
        return replace(self)


class QuadtreeNode:
    """Leaf (``domain`` set) or internal node with children in TL, TR, BL, BR order."""

    __slots__ = ("domain", "children")

    def __init__(self, domain: Optional[Domain], children: Optional[list["QuadtreeNode"]]) -> None:
        self.domain = domain
        self.children = children

    def is_leaf(self):
        return self.domain is not None

    @staticmethod
    def Leaf(domain: Domain) -> "QuadtreeNode":
        return QuadtreeNode(domain, None)

    @staticmethod
    def ParentOf(c1: "QuadtreeNode", c2: "QuadtreeNode", c3: "QuadtreeNode", c4: "QuadtreeNode") -> "QuadtreeNode":
        return QuadtreeNode(None, [c1, c2, c3, c4])


@dataclass
class QuadtreeImage:
    info: EncodingInfo
    forest: list[QuadtreeNode]


@dataclass
class ColoredQuadtreeImage:
    encoded_Y: QuadtreeImage
    encoded_Cb: QuadtreeImage
    encoded_Cr: QuadtreeImage


def get_dimension_bit_widths(info: EncodingInfo) -> list[int]:
    """Default (un-entropy-coded) field widths, indexed by CodingDimension.value (common.py:61-67)."""
    widths = [0] * len(CodingDimension)
    widths[CodingDimension.DOMAIN_ROW.value] = int(np.ceil(np.log2(info.img_height)))
    widths[CodingDimension.DOMAIN_COL.value] = int(np.ceil(np.log2(info.img_width)))
    widths[CodingDimension.SCALE.value] = info.scale_bits
    widths[CodingDimension.OFFSET.value] = info.offset_bits
    return widths


def get_domain_dimension_extractors() -> list[Callable[[Domain], int]]:
    ex: list = [None] * len(CodingDimension)
    ex[CodingDimension.DOMAIN_ROW.value] = lambda d: d.start_i
    ex[CodingDimension.DOMAIN_COL.value] = lambda d: d.start_j
    ex[CodingDimension.SCALE.value] = lambda d: d.quantized_scale
    ex[CodingDimension.OFFSET.value] = lambda d: d.quantized_offset
    return ex


# ---- flat stream <-> tree ------------------------------------------------------------------------
def stream_to_forest(nodes: np.ndarray, vals: np.ndarray, n_roots: int) -> list[QuadtreeNode]:
    """Rebuild the forest from the preorder stream of fic_qt_encode (kind, di, dj, orient, rot, qs, qo, exp)."""
    nodes = np.asarray(nodes)
    n = len(nodes)
    kind, di, dj, ori, rot, qs, qo = (nodes[:, c].tolist() for c in range(7)) if n else ([],) * 7
    sc, of = (vals[:, 0].tolist(), vals[:, 1].tolist()) if n else ([], [])
    gc_was_on = gc.isenabled()
    gc.disable()   # tens of thousands of small acyclic objects: the generational collector only gets in the way
    try:
        return _build_forest(kind, di, dj, ori, rot, qs, qo, sc, of, n, n_roots)
    finally:
        if gc_was_on:
            gc.enable()


def _build_forest(kind, di, dj, ori, rot, qs, qo, sc, of, n, n_roots):
    pos = 0
    forest: list[QuadtreeNode] = []
    new_node, new_domain = QuadtreeNode, Domain
    for _ in range(n_roots):
        # iterative preorder rebuild: stack of open internal nodes' child lists
        root = None
        stack: list[list] = []
        while True:
            if pos >= n:
                raise ValueError("node stream does not match the root count")
            if kind[pos] == 1:
                node = new_node(new_domain(di[pos], dj[pos], ori[pos], rot[pos], sc[pos], of[pos], qs[pos], qo[pos]), None)
                pos += 1
                if not stack:
                    root = node
                    break
                top = stack[-1]
                top.append(node)
                while len(top) == 4:
                    stack.pop()
                    if not stack:
                        break
                    top = stack[-1]
                if not stack:
                    break
            else:
                children: list = []
                node = new_node(None, children)
                pos += 1
                if stack:
                    stack[-1].append(node)
                else:
                    root = node
                stack.append(children)
        forest.append(root)
    if pos != n:
        raise ValueError("node stream does not match the root count")
    return forest


def iter_root_squares(width: int, height: int, i: int = 0, j: int = 0):
    """Yield (i, j, exponent) of the forest's roots in the reference's order: largest 2^k square, then
    the strip to its right, then the strip below (encoder.py:111-125, decoder.py:63-75)."""
    e = int(min(np.log2(width), np.log2(height)))
    side = 2 ** e
    yield i, j, e
    if side < width:
        yield from iter_root_squares(width - side, height, i, j + side)
    if side < height:
        yield from iter_root_squares(side, height - side, i + side, j)


def forest_to_leaves(qimg: QuadtreeImage, use_quantized_values: bool = True):
    """Flatten to the arrays fic_tree_decode takes: leaves[k] = (ri, rj, size, size, di, dj, orient, rot),
    so[k] = (scale, offset) de-quantised exactly as quadtree/decoder.py:90-98 does (the same float64 operations in
    the same order, applied to whole columns)."""
    info = qimg.info
    max_encoded_scale = np.float64(2 ** info.scale_bits)
    max_encoded_offset = np.float64(2 ** info.offset_bits)
    max_scale = info.max_scale
    leaves: list[tuple] = []
    doms: list[Domain] = []
    roots = list(iter_root_squares(info.img_width, info.img_height))
    if len(roots) != len(qimg.forest):
        raise ValueError(f"forest has {len(qimg.forest)} roots, the image geometry needs {len(roots)}")
    for (ri, rj, e), root in zip(roots, qimg.forest):
        stack = [(root, ri, rj, 2 ** e)]
        pop, push = stack.pop, stack.append
        while stack:
            node, i, j, size = pop()
            d = node.domain
            if d is not None:
                leaves.append((i, j, size, size, d.start_i, d.start_j, d.orientation, d.rotation))
                doms.append(d)
            else:
                half = size // 2
                c = node.children
                push((c[3], i + half, j + half, half))
                push((c[2], i + half, j, half))
                push((c[1], i, j + half, half))
                push((c[0], i, j, half))
    if use_quantized_values:
        qs = np.array([d.quantized_scale for d in doms], np.float64)
        qo = np.array([d.quantized_offset for d in doms], np.float64)
        scale = qs * (2 * max_scale) / max_encoded_scale - max_scale
        mul_coef = np.where(scale < 0, float(MAX_GRAY) * (1 + scale), float(MAX_GRAY))
        offset = qo / max_encoded_offset * mul_coef - MAX_GRAY * scale
        so = np.stack([scale, offset], axis=1) if len(doms) else np.zeros((0, 2))
    else:
        so = np.array([(float(d.scale), float(d.offset)) for d in doms], np.float64)
    return np.array(leaves, np.int32).reshape(-1, 8), np.ascontiguousarray(so, np.float64).reshape(-1, 2)
