"""HV file writer and reader (reference: hvv2/serialization.py:26-224).  Same container as `.fqt` with magic `hv`;
node stream breadth-first from the single root: leaf = `1`, row, col, orientation (1 bit), rotation // 2 (1 bit),
quantised scale, quantised offset; split = `0`, vertical (1 bit), split index (16 bits), then both children."""
from __future__ import annotations

import gc
from collections import deque
from typing import BinaryIO

import numpy as np

from _stream import BYTE_ORDER, open_target, patch_size, read_nodes_native, read_preamble, write_preamble
from bitbuffer import EOF, BitBuffer
from entropy.processor import CodingDimension, EntropyProcessor
from hvv2.common import Domain, HVImage, HVNode, SplitInfo
from quadtree.common import get_dimension_bit_widths, get_domain_dimension_extractors

HV_MAGIC = "hv".encode(encoding='ascii')
INTERNAL_NODE_BIT = 0
LEAF_BIT = 1
STRUCT_BYTE_ORDER = BYTE_ORDER
HEADER_SIZE = len(HV_MAGIC) + 14
UNSUPPORTED = 0.


class HVSerializer:
    def __init__(self, entropy_processor: EntropyProcessor, entropy_coding_dimensions: set[CodingDimension],
                 required_entropy_coding_gain=1) -> None:
        self.entropy_processor = entropy_processor
        self.entropy_coding_dimensions = entropy_coding_dimensions
        self.required_entropy_coding_gain = required_entropy_coding_gain
        self.file: BinaryIO = None

    def serialize(self, hv_img: HVImage, output: str | BinaryIO):
        self.img = hv_img
        gc_was_on = gc.isenabled()
        gc.disable()   # tens of thousands of short-lived tuples and nodes, none of them cyclic
        try:
            self._walk()
            self._choose_coded_dimensions()
            with open_target(output, 'wb') as f:
                self.file = f
                start = f.tell()
                bits = BitBuffer(f, BYTE_ORDER)
                write_preamble(f, bits, HV_MAGIC, hv_img.info, self.entropy_processor, self.entropy_coding_dimensions)
                self._write_nodes(bits)
                patch_size(f, start, HV_MAGIC)
        finally:
            if gc_was_on:
                gc.enable()

    def _walk(self):
        """One breadth-first pass over the tree (the order of the node stream and of the Huffman statistics alike):
        per node its kind and raw fields, as columns."""
        flat = self.img.flat_stream()
        if flat is not None:
            # nodes of one level appear in a preorder stream from left to right: breadth-first = stable sort by depth
            import fic_b200
            nodes = flat[0]
            info = self.img.info
            _, depth = fic_b200.hv_layout(nodes, info.img_height, info.img_width)
            table = nodes[np.argsort(depth, kind="stable")][:, :7].astype(np.int64)
            leaf = table[:, 0] == LEAF_BIT
            table[leaf, 4] //= 2                       # rotation is stored as rotation // 2 (one bit)
            self._kind = table[:, 0]
            self._fields = [table[:, c] for c in range(1, 7)]
            return
        rows = []
        add = rows.append
        pending = deque([self.img.root])
        pop, push = pending.popleft, pending.append
        while pending:
            node = pop()
            d = node.domain
            if d is not None:
                add((LEAF_BIT, d.start_i, d.start_j, d.orientation, d.rotation // 2, d.quantized_scale, d.quantized_offset))
            else:
                si = node.split_info
                add((INTERNAL_NODE_BIT, int(si.vertical_split), si.split_idx, 0, 0, 0, 0))
                push(node.c1)
                push(node.c2)
        table = np.array(rows, np.int64).reshape(-1, 7)
        self._kind = table[:, 0]
        self._fields = [table[:, c] for c in range(1, 7)]

    def _choose_coded_dimensions(self):
        values: dict[CodingDimension, list[int]] = {}
        if self.entropy_coding_dimensions:
            leaf = self._kind == LEAF_BIT
            column = {CodingDimension.DOMAIN_ROW: 0, CodingDimension.DOMAIN_COL: 1, CodingDimension.SCALE: 4,
                      CodingDimension.OFFSET: 5}
            values = {dim: np.ascontiguousarray(self._fields[column[dim]][leaf]) for dim in CodingDimension
                      if dim in self.entropy_coding_dimensions}
        gains = self.entropy_processor.prepare_for_encoding(values, get_dimension_bit_widths(self.img.info))
        self.entropy_coding_dimensions = {dim for dim, gain in gains.items() if gain >= self.required_entropy_coding_gain}

    def _coded(self, column: np.ndarray, dim: CodingDimension):
        """(code, width) of every value of a column through the entropy processor."""
        return self.entropy_processor.encode_column(column, dim)

    def _write_nodes(self, bits: BitBuffer):
        # leaf = 1, row, col, orientation (1), rotation // 2 (1), scale, offset; split = 0, vertical (1), index (16)
        kind, f = self._kind, self._fields
        leaf = kind == LEAF_BIT
        n = kind.size
        vals = np.zeros((n, 7), np.uint64)
        wid = np.zeros((n, 7), np.int64)
        vals[:, 0] = kind
        wid[:, 0] = 1
        row_c, row_w = self._coded(f[0][leaf], CodingDimension.DOMAIN_ROW)
        col_c, col_w = self._coded(f[1][leaf], CodingDimension.DOMAIN_COL)
        sc_c, sc_w = self._coded(f[4][leaf], CodingDimension.SCALE)
        of_c, of_w = self._coded(f[5][leaf], CodingDimension.OFFSET)
        vals[leaf, 1], wid[leaf, 1] = row_c, row_w
        vals[leaf, 2], wid[leaf, 2] = col_c, col_w
        vals[leaf, 3], wid[leaf, 3] = f[2][leaf], 1
        vals[leaf, 4], wid[leaf, 4] = f[3][leaf], 1
        vals[leaf, 5], wid[leaf, 5] = sc_c, sc_w
        vals[leaf, 6], wid[leaf, 6] = of_c, of_w
        split = ~leaf
        vals[split, 1], wid[split, 1] = f[0][split], 1
        vals[split, 2], wid[split, 2] = f[1][split], 16
        bits.write_many(vals.ravel(), wid.ravel())      # zero-width slots vanish
        if not bits.is_empty():
            bits.flush()


class HVDeserializer:
    def __init__(self) -> None:
        self.file: BinaryIO = None
        self.img: HVImage = None
        self.entropy_processor: EntropyProcessor = None
        self.size = 0

    def deserialize(self, input: str | BinaryIO) -> HVImage:
        with open_target(input, 'rb') as f:
            self.file = f
            bits = BitBuffer(f, BYTE_ORDER)
            info, self.size, self.entropy_processor = read_preamble(f, bits, HV_MAGIC, 'hv')
            nodes = read_nodes_native(f, bits, self.entropy_processor, fic_tree_kind=1, root_exps=None,
                                      cap=2 * info.img_width * info.img_height + 64)
            if nodes is not None:
                self.img = HVImage(info, stream=(nodes, np.zeros((nodes.shape[0], 2))))
            else:
                self.img = HVImage(info, self._read_tree(bits))
            return self.img

    def _read_tree(self, bits: BitBuffer) -> HVNode:
        dec = self.entropy_processor.decode
        root = None
        waiting = deque()              # (parent, which child) in stream order
        while root is None or waiting:
            kind = bits.read(1)
            assert kind != EOF, "Invalid size encoding, buffer returned EOF"
            if kind == LEAF_BIT:
                start_i = dec(bits, CodingDimension.DOMAIN_ROW)
                start_j = dec(bits, CodingDimension.DOMAIN_COL)
                orientation = bits.read(1)
                rotation = bits.read(1) * 2
                qs = dec(bits, CodingDimension.SCALE)
                qo = dec(bits, CodingDimension.OFFSET)
                node = HVNode.Leaf(Domain(start_i, start_j, orientation, rotation, scale=UNSUPPORTED, offset=UNSUPPORTED,
                                          quantized_scale=qs, quantized_offset=qo))
            else:
                vertical = bool(bits.read(1))
                node = HVNode.ParentOf(None, None, SplitInfo(vertical, bits.read(16)))
            if root is None:
                root = node
            else:
                parent, which = waiting.popleft()
                if which == 0:
                    parent.c1 = node
                else:
                    parent.c2 = node
            if not node.is_leaf():
                waiting.append((node, 0))
                waiting.append((node, 1))
        return root
