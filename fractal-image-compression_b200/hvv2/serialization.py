"""HV file writer and reader (reference: hvv2/serialization.py:26-224).  Same container as `.fqt` with magic `hv`;
node stream breadth-first from the single root: leaf = `1`, row, col, orientation (1 bit), rotation // 2 (1 bit),
quantised scale, quantised offset; split = `0`, vertical (1 bit), split index (16 bits), then both children."""
from __future__ import annotations

from collections import deque
from typing import BinaryIO

from _stream import BYTE_ORDER, open_target, patch_size, read_preamble, write_preamble
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
        self._choose_coded_dimensions()
        with open_target(output, 'wb') as f:
            self.file = f
            start = f.tell()
            bits = BitBuffer(f, BYTE_ORDER)
            write_preamble(f, bits, HV_MAGIC, hv_img.info, self.entropy_processor, self.entropy_coding_dimensions)
            self._write_nodes(bits)
            patch_size(f, start, HV_MAGIC)

    def _choose_coded_dimensions(self):
        values: dict[CodingDimension, list[int]] = {}
        if self.entropy_coding_dimensions:
            extract = get_domain_dimension_extractors()
            wanted = [dim for dim in CodingDimension if dim in self.entropy_coding_dimensions]
            values = {dim: [] for dim in wanted}
            pending = deque([self.img.root])
            while pending:
                node = pending.popleft()
                if node.is_leaf():
                    for dim in wanted:
                        values[dim].append(extract[dim.value](node.domain))
                else:
                    pending.append(node.c1)
                    pending.append(node.c2)
        gains = self.entropy_processor.prepare_for_encoding(values, get_dimension_bit_widths(self.img.info))
        self.entropy_coding_dimensions = {dim for dim, gain in gains.items() if gain >= self.required_entropy_coding_gain}

    def _write_nodes(self, bits: BitBuffer):
        enc = self.entropy_processor.encode
        pending = deque([self.img.root])
        while pending:
            node = pending.popleft()
            if node.is_leaf():
                d = node.domain
                bits.write(LEAF_BIT, 1)
                bits.write(*enc(d.start_i, CodingDimension.DOMAIN_ROW))
                bits.write(*enc(d.start_j, CodingDimension.DOMAIN_COL))
                bits.write(d.orientation, 1)
                bits.write(d.rotation // 2, 1)
                bits.write(*enc(d.quantized_scale, CodingDimension.SCALE))
                bits.write(*enc(d.quantized_offset, CodingDimension.OFFSET))
            else:
                bits.write(INTERNAL_NODE_BIT, 1)
                bits.write(int(node.split_info.vertical_split), 1)
                bits.write(node.split_info.split_idx, 16)
                pending.append(node.c1)
                pending.append(node.c2)
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
