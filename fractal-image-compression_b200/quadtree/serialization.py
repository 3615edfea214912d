"""`.fqt` writer and reader (reference: quadtree/serialization.py:26-223), byte-compatible with the reference's
files: the two bundled fixtures round-trip identically (tests/test_serialization.py).

Node stream, breadth-first per root: leaf = `1`, row, col, orientation (1 bit), rotation (2 bits), quantised scale,
quantised offset; internal node = `0` followed (later in the stream) by its four children.
"""
from __future__ import annotations

import gc
from collections import deque
from typing import BinaryIO

import numpy as np

from _stream import BYTE_ORDER, open_target, patch_size, read_nodes_native, read_preamble, write_preamble
from bitbuffer import EOF, BitBuffer
from entropy.processor import CodingDimension, EntropyProcessor
from quadtree.common import (Domain, QuadtreeImage, QuadtreeNode, get_dimension_bit_widths,
                             get_domain_dimension_extractors, iter_root_squares, stream_geometry)

QUADTREE_MAGIC = "quadtree".encode(encoding='ascii')
INTERNAL_NODE_BIT = 0
LEAF_BIT = 1
STRUCT_BYTE_ORDER = BYTE_ORDER
HEADER_SIZE = len(QUADTREE_MAGIC) + 14
UNSUPPORTED = 0.


def _leaves_breadth_first(forest):
    """Leaf domains in the order the reference collects them for the Huffman statistics: ONE queue over all roots
    (serialization.py:62-78), which differs from the per-root order of the node stream."""
    pending = deque(forest)
    while pending:
        node = pending.popleft()
        if node.is_leaf():
            yield node.domain
        else:
            pending.extend(node.children)


class QuadtreeSerializer:
    def __init__(self, entropy_processor: EntropyProcessor, entropy_coding_dimensions: set[CodingDimension],
                 required_entropy_coding_gain=1) -> None:
        self.entropy_processor = entropy_processor
        self.entropy_coding_dimensions = entropy_coding_dimensions
        self.required_entropy_coding_gain = required_entropy_coding_gain
        self.file: BinaryIO = None

    def serialize(self, quadtree_img: QuadtreeImage, output: str | BinaryIO):
        self.img = quadtree_img
        gc_was_on = gc.isenabled()
        gc.disable()   # tens of thousands of short-lived tuples and nodes, none of them cyclic
        try:
            self._choose_coded_dimensions()
            with open_target(output, 'wb') as f:
                self.file = f
                start = f.tell()
                bits = BitBuffer(f, BYTE_ORDER)
                write_preamble(f, bits, QUADTREE_MAGIC, quadtree_img.info, self.entropy_processor, self.entropy_coding_dimensions)
                self._write_nodes(bits)
                patch_size(f, start, QUADTREE_MAGIC)
        finally:
            if gc_was_on:
                gc.enable()

    def _choose_coded_dimensions(self):
        values: dict[CodingDimension, list[int]] = {}
        flat = self.img.flat_stream()
        self._geometry = None
        if flat is not None:
            info = self.img.info
            self._geometry = stream_geometry(flat[0], flat[2], info.img_width, info.img_height)
        if self.entropy_coding_dimensions:
            if flat is not None:
                # the single queue over all roots = level by level (relative to each root), roots in order, Morton inside
                nodes = flat[0]
                root_of, _, _, e, off = self._geometry
                leaf = np.flatnonzero(nodes[:, 0] == LEAF_BIT)
                depth = e[np.searchsorted(root_of, root_of[leaf])] - e[leaf]     # a root's node is its first in the stream
                order = leaf[np.lexsort((off[leaf], root_of[leaf], depth))]
                table = nodes[order][:, [1, 2, 5, 6]].astype(np.int64)
            else:
                rows = [(d.start_i, d.start_j, d.quantized_scale, d.quantized_offset)
                        for d in _leaves_breadth_first(self.img.forest)]
                table = np.array(rows, np.int64).reshape(-1, 4)
            column = {CodingDimension.DOMAIN_ROW: 0, CodingDimension.DOMAIN_COL: 1, CodingDimension.SCALE: 2,
                      CodingDimension.OFFSET: 3}
            values = {dim: np.ascontiguousarray(table[:, column[dim]]) for dim in CodingDimension if dim in self.entropy_coding_dimensions}
        gains = self.entropy_processor.prepare_for_encoding(values, get_dimension_bit_widths(self.img.info))
        # the reference overwrites the requested set with the profitable one (serialization.py:60)
        self.entropy_coding_dimensions = {dim for dim, gain in gains.items() if gain >= self.required_entropy_coding_gain}

    def _coded(self, column: np.ndarray, dim: CodingDimension):
        """(code, width) of every value of a column through the entropy processor."""
        return self.entropy_processor.encode_column(column, dim)

    def _write_nodes(self, bits: BitBuffer):
        # node stream, breadth-first per root: leaf = 1, row, col, orientation (1), rotation (2), scale, offset; inner = 0
        flat = self.img.flat_stream()
        if flat is not None and self._geometry is not None:
            # preorder stream -> breadth first per root: by root, then level (larger blocks first), then Morton offset
            root_of, _, _, e, off = self._geometry
            table = flat[0][np.lexsort((off, -e, root_of))][:, :7].astype(np.int64)
        else:
            rows = []
            add = rows.append
            for root in self.img.forest:
                pending = deque([root])
                pop = pending.popleft
                while pending:
                    node = pop()
                    d = node.domain
                    if d is not None:
                        add((LEAF_BIT, d.start_i, d.start_j, d.orientation, d.rotation, d.quantized_scale, d.quantized_offset))
                    else:
                        add((INTERNAL_NODE_BIT, 0, 0, 0, 0, 0, 0))
                        pending.extend(node.children)
            table = np.array(rows, np.int64).reshape(-1, 7)
        leaf = table[:, 0] == LEAF_BIT
        n = table.shape[0]
        vals = np.zeros((n, 7), np.uint64)
        wid = np.zeros((n, 7), np.int64)
        vals[:, 0] = table[:, 0]
        wid[:, 0] = 1
        for col, dim in ((1, CodingDimension.DOMAIN_ROW), (2, CodingDimension.DOMAIN_COL), (5, CodingDimension.SCALE),
                         (6, CodingDimension.OFFSET)):
            vals[leaf, col], wid[leaf, col] = self._coded(table[leaf, col], dim)
        vals[leaf, 3], wid[leaf, 3] = table[leaf, 3], 1
        vals[leaf, 4], wid[leaf, 4] = table[leaf, 4], 2
        bits.write_many(vals.ravel(), wid.ravel())      # zero-width slots vanish
        if not bits.is_empty():
            bits.flush()


class QuadtreeDeserializer:
    def __init__(self) -> None:
        self.file: BinaryIO = None
        self.img: QuadtreeImage = None
        self.entropy_processor: EntropyProcessor = None
        self.size = 0

    def deserialize(self, input: str | BinaryIO) -> QuadtreeImage:
        with open_target(input, 'rb') as f:
            self.file = f
            bits = BitBuffer(f, BYTE_ORDER)
            info, self.size, self.entropy_processor = read_preamble(f, bits, QUADTREE_MAGIC, 'quadtree')
            exps = [e for _, _, e in iter_root_squares(info.img_width, info.img_height)]
            nodes = read_nodes_native(f, bits, self.entropy_processor, fic_tree_kind=0, root_exps=exps,
                                      cap=2 * info.img_width * info.img_height + 64)
            if nodes is not None:
                self.img = QuadtreeImage(info, stream=(nodes, np.zeros((nodes.shape[0], 3)), len(exps)))
            else:
                self.img = QuadtreeImage(info, self._read_forest(bits, info))
            return self.img

    def _read_forest(self, bits: BitBuffer, info) -> list[QuadtreeNode]:
        dec = self.entropy_processor.decode
        forest = []
        # the image geometry fixes the number of roots (encoder.py:111-125); each root is a breadth-first stream
        for _ in iter_root_squares(info.img_width, info.img_height):
            root = None
            waiting = deque()          # (parent, child index) slots still to be filled, in stream order
            while root is None or waiting:
                kind = bits.read(1)
                assert kind != EOF, "Invalid size encoding, buffer returned EOF"
                if kind == LEAF_BIT:
                    start_i = dec(bits, CodingDimension.DOMAIN_ROW)
                    start_j = dec(bits, CodingDimension.DOMAIN_COL)
                    orientation = bits.read(1)
                    rotation = bits.read(2)
                    qs = dec(bits, CodingDimension.SCALE)
                    qo = dec(bits, CodingDimension.OFFSET)
                    node = QuadtreeNode.Leaf(Domain(start_i, start_j, orientation, rotation, scale=UNSUPPORTED,
                                                    offset=UNSUPPORTED, quantized_scale=qs, quantized_offset=qo))
                else:
                    node = QuadtreeNode.ParentOf(None, None, None, None)
                if root is None:
                    root = node
                else:
                    parent, k = waiting.popleft()
                    parent.children[k] = node
                if not node.is_leaf():
                    waiting.extend((node, k) for k in range(4))
            forest.append(root)
        return forest
