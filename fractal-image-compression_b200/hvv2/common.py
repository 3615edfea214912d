"""Data model of the HV codec (reference: hvv2/common.py:7-45) plus the flat-array converters used by
the C ABI (include/fic.h: fic_hv_encode, fic_tree_decode)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import numpy as np

from quadtree.common import MAX_GRAY, EncodingInfo


@dataclass
class Domain:
    start_i: int
    start_j: int
    orientation: int
    rotation: int
    scale: float
    offset: float
    quantized_scale: int
    quantized_offset: int


@dataclass
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


@dataclass
class HVImage:
    info: EncodingInfo
    root: HVNode


def stream_to_tree(nodes: np.ndarray, vals: np.ndarray) -> HVNode:
    """Rebuild the binary tree from the preorder stream of fic_hv_encode."""
    rows = nodes.tolist()
    vv = vals.tolist()
    root = None
    stack: list[list] = []  # [node, children attached]
    for r, v in zip(rows, vv):
        if r[0] == 1:
            node = HVNode.Leaf(Domain(r[1], r[2], r[3], r[4], v[0], v[1], r[5], r[6]))
        else:
            node = HVNode(None, None, SplitInfo(bool(r[1]), int(r[2])), None)
        if stack:
            top = stack[-1]
            if top[1] == 0:
                top[0].c1 = node
            else:
                top[0].c2 = node
            top[1] += 1
        else:
            root = node
        if not node.is_leaf():
            stack.append([node, 0])
        while stack and stack[-1][1] == 2:
            stack.pop()
    if stack or root is None:
        raise ValueError("truncated HV node stream")
    return root


def tree_to_leaves(hv_img: HVImage, use_quantized_values: bool = True):
    """leaves[k] = (ri, rj, h, w, di, dj, orient, rot), so[k] = (scale, offset) as hvv2/decoder.py:59-74."""
    info = hv_img.info
    max_encoded_scale = np.float64(2 ** info.scale_bits)
    max_encoded_offset = np.float64(2 ** info.offset_bits)
    max_scale = info.max_scale
    leaves, so = [], []
    stack = [(hv_img.root, 0, 0, info.img_width, info.img_height)]
    while stack:
        node, i, j, w, h = stack.pop()
        if node.is_leaf():
            d = node.domain
            if use_quantized_values:
                scale = np.float64(d.quantized_scale) * (2 * max_scale) / max_encoded_scale - max_scale
                mul_coef = float(MAX_GRAY)
                if scale < 0:
                    mul_coef *= (1 + scale)
                offset = np.float64(d.quantized_offset) / max_encoded_offset * mul_coef - MAX_GRAY * scale
            else:
                scale, offset = d.scale, d.offset
            leaves.append((i, j, h, w, d.start_i, d.start_j, d.orientation, d.rotation))
            so.append((float(scale), float(offset)))
        else:
            s = node.split_info.split_idx
            if node.split_info.vertical_split:
                stack.append((node.c2, i, j + s + 1, w - s - 1, h))
                stack.append((node.c1, i, j, s + 1, h))
            else:
                stack.append((node.c2, i + s + 1, j, w, h - s - 1))
                stack.append((node.c1, i, j, w, s + 1))
    return np.array(leaves, np.int32).reshape(-1, 8), np.array(so, np.float64).reshape(-1, 2)
