"""Data model of the HV codec (reference: hvv2/common.py:7-45) plus the flat-array converters used by
the C ABI (include/fic.h: fic_hv_encode, fic_tree_decode)."""
from __future__ import annotations

import gc
from dataclasses import dataclass
from typing import Optional

import numpy as np

from quadtree.common import MAX_GRAY, EncodingInfo


@dataclass(slots=True)
class Domain:
    start_i: int
    start_j: int
    orientation: int
    rotation: int
    scale: float
    offset: float
    quantized_scale: int
    quantized_offset: int


@dataclass(slots=True)
class SplitInfo:
    vertical_split: bool
    split_idx: int


class HVNode:
    """Leaf (``domain`` set) or a split: ``c1`` is the left/top part of size ``split_idx + 1``."""

    __slots__ = ("c1", "c2", "split_info", "domain")

    def __init__(self, c1: Optional["HVNode"], c2: Optional["HVNode"], split_info: Optional[SplitInfo],
                 domain: Optional[Domain]) -> None:
        self.c1 = c1
        self.c2 = c2
        self.split_info = split_info
        self.domain = domain

    def is_leaf(self) -> bool:
        return self.domain is not None

    @staticmethod
    def ParentOf(c1: "HVNode", c2: "HVNode", split_info: SplitInfo) -> "HVNode":
        return HVNode(c1, c2, split_info, None)

    @staticmethod
    def Leaf(domain: Domain) -> "HVNode":
        return HVNode(None, None, None, domain)


class HVImage:
    """`info` + `root`, as in the reference (hvv2/common.py:42-45).  An image made by this package's encoder or reader
    also carries the flat preorder stream; the `HVNode` objects are built when `root` is first read, and from then on the
    stream is dropped (the objects may be edited).  Until then decoder and writer work on the stream (`flat_stream()`)."""

    __slots__ = ("info", "_root", "_stream")

    def __init__(self, info: EncodingInfo, root: Optional[HVNode] = None, *, stream=None) -> None:
        if root is None and stream is None:
            raise TypeError("HVImage needs a root")
        self.info = info
        self._root = root
        self._stream = stream if root is None else None

    @property
    def root(self) -> HVNode:
        if self._root is None:
            self._root = stream_to_tree(*self._stream)
            self._stream = None
        return self._root

    @root.setter
    def root(self, value: HVNode) -> None:
        self._root = value
        self._stream = None

    def flat_stream(self):
        """(nodes[k,8] int32, vals[k,>=2] float64) while nobody has seen the node objects, else None."""
        return self._stream

    def __eq__(self, other) -> bool:
        return isinstance(other, HVImage) and self.info == other.info and self.root == other.root

    def __repr__(self) -> str:
        return f"HVImage(info={self.info!r})"


def stream_to_tree(nodes: np.ndarray, vals: np.ndarray) -> HVNode:
    """Rebuild the binary tree from the preorder stream of fic_hv_encode."""
    nodes = np.asarray(nodes)
    n = len(nodes)
    if n == 0:
        raise ValueError("truncated HV node stream")
    cols = [nodes[:, c].tolist() for c in range(7)]
    sc, of = vals[:, 0].tolist(), vals[:, 1].tolist()
    gc_was_on = gc.isenabled()
    gc.disable()   # tens of thousands of small acyclic objects: the generational collector only gets in the way
    try:
        return _build_tree(cols, sc, of, n)
    finally:
        if gc_was_on:
            gc.enable()


def _build_tree(cols, sc, of, n) -> HVNode:
    kind, c1, c2, c3, c4, c5, c6 = cols
    root = None
    stack: list[HVNode] = []   # open splits: c1 is None until the first child arrives, c2 until the second
    new_node, new_domain, new_split = HVNode, Domain, SplitInfo
    for k in range(n):
        if root is not None and not stack:
            raise ValueError("HV node stream continues after the tree is complete")
        if kind[k] == 1:
            node = new_node(None, None, None, new_domain(c1[k], c2[k], c3[k], c4[k], sc[k], of[k], c5[k], c6[k]))
        else:
            node = new_node(None, None, new_split(bool(c1[k]), c2[k]), None)
        if stack:
            top = stack[-1]
            if top.c1 is None:
                top.c1 = node
            else:
                top.c2 = node
                stack.pop()
        else:
            root = node
        if kind[k] != 1:
            stack.append(node)
    if stack or root is None:
        raise ValueError("truncated HV node stream")
    return root


def _dequantise(qs: np.ndarray, qo: np.ndarray, info: EncodingInfo) -> np.ndarray:
    """hvv2/decoder.py:64-71 on whole columns: the same float64 operations in the same order."""
    max_encoded_scale = np.float64(2 ** info.scale_bits)
    max_encoded_offset = np.float64(2 ** info.offset_bits)
    max_scale = info.max_scale
    scale = qs * (2 * max_scale) / max_encoded_scale - max_scale
    mul_coef = np.where(scale < 0, float(MAX_GRAY) * (1 + scale), float(MAX_GRAY))
    offset = qo / max_encoded_offset * mul_coef - MAX_GRAY * scale
    return np.stack([scale, offset], axis=1) if len(qs) else np.zeros((0, 2))


def tree_to_leaves(hv_img: HVImage, use_quantized_values: bool = True):
    """leaves[k] = (ri, rj, h, w, di, dj, orient, rot), so[k] = (scale, offset) as hvv2/decoder.py:59-88 (the same float64
    operations in the same order, applied to whole columns).  Leaves come out in preorder."""
    info = hv_img.info
    flat = hv_img.flat_stream()
    if flat is not None:
        import fic_b200
        nodes, vals = flat
        rects, _ = fic_b200.hv_layout(nodes, info.img_height, info.img_width)
        leaf = nodes[:, 0] == 1
        leaves = np.empty((int(leaf.sum()), 8), np.int32)
        leaves[:, 0], leaves[:, 1], leaves[:, 2], leaves[:, 3] = rects[leaf, 0], rects[leaf, 1], rects[leaf, 3], rects[leaf, 2]
        leaves[:, 4:8] = nodes[leaf, 1:5]
        if use_quantized_values:
            so = _dequantise(nodes[leaf, 5].astype(np.float64), nodes[leaf, 6].astype(np.float64), info)
        else:
            so = vals[leaf, :2]
        return leaves, np.ascontiguousarray(so, np.float64).reshape(-1, 2)
    leaves, doms = [], []
    stack = [(hv_img.root, 0, 0, info.img_width, info.img_height)]
    pop, push = stack.pop, stack.append
    while stack:
        node, i, j, w, h = pop()
        d = node.domain
        if d is not None:
            leaves.append((i, j, h, w, d.start_i, d.start_j, d.orientation, d.rotation))
            doms.append(d)
        else:
            s = node.split_info.split_idx
            if node.split_info.vertical_split:
                push((node.c2, i, j + s + 1, w - s - 1, h))
                push((node.c1, i, j, s + 1, h))
            else:
                push((node.c2, i + s + 1, j, w, h - s - 1))
                push((node.c1, i, j, w, s + 1))
    if use_quantized_values:
        so = _dequantise(np.array([d.quantized_scale for d in doms], np.float64),
                         np.array([d.quantized_offset for d in doms], np.float64), info)
    else:
        so = np.array([(float(d.scale), float(d.offset)) for d in doms], np.float64)
    return np.array(leaves, np.int32).reshape(-1, 8), np.ascontiguousarray(so, np.float64).reshape(-1, 2)
