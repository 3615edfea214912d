"""Data model of the quadtree codec (reference: quadtree/common.py:10-75).

Same class and field names as the reference so its serialisers, decoders and user code work on
these objects unchanged.  ``stream_to_forest`` / ``forest_to_leaves`` convert between the tree
objects and the flat arrays of the C ABI (include/fic.h: fic_qt_encode, fic_tree_decode).
"""
from __future__ import annotations

import functools
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


class QuadtreeImage:
    """`info` + `forest`, as in the reference (common.py:45-48).

    An image made by this package's encoder also carries the flat preorder stream `fic_qt_encode` returned.  The
    `QuadtreeNode` / `Domain` objects of `forest` are created when somebody first asks for them -- and from that moment
    the stream is dropped, because the objects may be edited.  Until then the decoder and the serialiser of this package
    work on the stream with array operations (`flat_stream()`), which is what keeps encode -> decode and encode -> file at
    kernel speed instead of at the speed of building and walking ~10^5 Python objects."""

    __slots__ = ("info", "_forest", "_stream")

    def __init__(self, info: EncodingInfo, forest: Optional[list[QuadtreeNode]] = None, *, stream=None) -> None:
        if forest is None and stream is None:
            raise TypeError("QuadtreeImage needs a forest")
        self.info = info
        self._forest = forest
        self._stream = stream if forest is None else None

    @property
    def forest(self) -> list[QuadtreeNode]:
        if self._forest is None:
            self._forest = stream_to_forest(*self._stream)
            self._stream = None
        return self._forest

    @forest.setter
    def forest(self, value: list[QuadtreeNode]) -> None:
        self._forest = value
        self._stream = None

    def flat_stream(self):
        """(nodes[k,8] int32, vals[k,3] float64, n_roots) while nobody has seen the node objects, else None."""
        return self._stream

    def __eq__(self, other) -> bool:
        return isinstance(other, QuadtreeImage) and self.info == other.info and self.forest == other.forest

    def __repr__(self) -> str:
        return f"QuadtreeImage(info={self.info!r}, forest=<{len(self.forest)} roots>)"


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


@functools.lru_cache(maxsize=64)
def _root_squares(width: int, height: int) -> tuple:
    """(i, j, exponent) of every root for a width x height image, in the reference's order."""
    out = []
    stack = [(0, 0, width, height)]
    while stack:                                   # preorder: the square, then the strip to its right, then the one below
        i, j, w, h = stack.pop()
        e = min(w.bit_length(), h.bit_length()) - 1   # int(min(log2 w, log2 h)) without float rounding
        side = 1 << e
        out.append((i, j, e))
        if side < h:
            stack.append((i + side, j, side, h - side))
        if side < w:
            stack.append((i, j + side, w - side, h))
    return tuple(out)


def iter_root_squares(width: int, height: int, i: int = 0, j: int = 0):
    """Yield (i, j, exponent) of the forest's roots in the reference's order: largest 2^k square, then
    the strip to its right, then the strip below (encoder.py:111-125, decoder.py:63-75).  A ragged image has hundreds of
    roots (313 x 413: 950), so the list is built iteratively and cached per image size."""
    for ri, rj, e in _root_squares(int(width), int(height)):
        yield ri + i, rj + j, e


def _compact_even_bits(v: np.ndarray) -> np.ndarray:
    """Bits 0, 2, 4, ... of v packed together (inverse of the Morton interleave), v < 2^32."""
    v = v & np.uint64(0x55555555)
    v = (v | (v >> np.uint64(1))) & np.uint64(0x33333333)
    v = (v | (v >> np.uint64(2))) & np.uint64(0x0f0f0f0f)
    v = (v | (v >> np.uint64(4))) & np.uint64(0x00ff00ff)
    v = (v | (v >> np.uint64(8))) & np.uint64(0x0000ffff)
    return v


def stream_geometry(nodes: np.ndarray, n_roots: int, width: int, height: int):
    """Where every node of a preorder stream sits: (root index, row, col, exponent, offset) per node, as int64 arrays.

    Children are emitted TL, TR, BL, BR, so inside a root the leaves appear in Morton order and a node starts at the
    Morton offset equal to the leaf area emitted before it; de-interleaving that offset gives its corner.  `offset` is
    that Morton offset inside the root (breadth-first order of one level = ascending offset)."""
    nodes = np.asarray(nodes)
    roots = list(iter_root_squares(width, height))
    if len(roots) != n_roots:
        raise ValueError(f"forest has {n_roots} roots, the image geometry needs {len(roots)}")
    e = nodes[:, 7].astype(np.int64)
    area = np.where(nodes[:, 0] == 1, np.int64(1) << (2 * e), np.int64(0))
    before = np.cumsum(area) - area
    root_area = np.array([np.int64(1) << (2 * re) for _, _, re in roots], np.int64)
    root_start = np.cumsum(root_area) - root_area
    if int(area.sum()) != int(root_area.sum()):
        raise ValueError("node stream does not cover the image")
    root_of = np.searchsorted(root_start, before, side='right') - 1
    off = before - root_start[root_of]
    u = off.astype(np.uint64)
    row = np.array([r[0] for r in roots], np.int64)[root_of] + _compact_even_bits(u >> np.uint64(1)).astype(np.int64)
    col = np.array([r[1] for r in roots], np.int64)[root_of] + _compact_even_bits(u).astype(np.int64)
    return root_of, row, col, e, off


def _dequantise(qs: np.ndarray, qo: np.ndarray, info: EncodingInfo) -> np.ndarray:
    """quadtree/decoder.py:90-98 on whole columns: the same float64 operations in the same order."""
    max_encoded_scale = np.float64(2 ** info.scale_bits)
    max_encoded_offset = np.float64(2 ** info.offset_bits)
    max_scale = info.max_scale
    scale = qs * (2 * max_scale) / max_encoded_scale - max_scale
    mul_coef = np.where(scale < 0, float(MAX_GRAY) * (1 + scale), float(MAX_GRAY))
    offset = qo / max_encoded_offset * mul_coef - MAX_GRAY * scale
    return np.stack([scale, offset], axis=1) if len(qs) else np.zeros((0, 2))


def forest_to_leaves(qimg: QuadtreeImage, use_quantized_values: bool = True):
    """Flatten to the arrays fic_tree_decode takes: leaves[k] = (ri, rj, size, size, di, dj, orient, rot),
    so[k] = (scale, offset) de-quantised exactly as quadtree/decoder.py:90-98 does (the same float64 operations in
    the same order, applied to whole columns).  Leaves come out in preorder."""
    info = qimg.info
    flat = qimg.flat_stream()
    if flat is not None:
        nodes, vals, n_roots = flat
        _, row, col, e, _ = stream_geometry(nodes, n_roots, info.img_width, info.img_height)
        leaf = nodes[:, 0] == 1
        side = (np.int64(1) << e[leaf])
        leaves = np.empty((int(leaf.sum()), 8), np.int32)
        leaves[:, 0], leaves[:, 1], leaves[:, 2], leaves[:, 3] = row[leaf], col[leaf], side, side
        leaves[:, 4:8] = nodes[leaf, 1:5]
        if use_quantized_values:
            so = _dequantise(nodes[leaf, 5].astype(np.float64), nodes[leaf, 6].astype(np.float64), info)
        else:
            so = vals[leaf, :2]
        return leaves, np.ascontiguousarray(so, np.float64).reshape(-1, 2)
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
        so = _dequantise(np.array([d.quantized_scale for d in doms], np.float64),
                         np.array([d.quantized_offset for d in doms], np.float64), info)
    else:
        so = np.array([(float(d.scale), float(d.offset)) for d in doms], np.float64)
    return np.array(leaves, np.int32).reshape(-1, 8), np.ascontiguousarray(so, np.float64).reshape(-1, 2)
