"""Shared container logic of the `.fqt` (quadtree) and HV files: header, entropy-coder preamble, size patch.

Layout (reference quadtree/serialization.py:13-101, hvv2/serialization.py:13-103):
  header  struct '!{magic}sfBBHHI' = magic, max_scale f32, scale_bits u8, offset_bits u8, width u16, height u16,
          total size u32 (patched after the payload is written)
  bits    8 bits entropy-processor id; 4 presence bits in CodingDimension order; one breadth-first Huffman tree per
          present dimension; then the node stream; zero padded to a 32-bit word.
"""
from __future__ import annotations

import struct
from typing import BinaryIO

from bitbuffer import BitBuffer
from entropy.processor import (ENTROPY_PROCESSOR_ID_BITS, CodingDimension, EntropyProcessor, EntropyProcessorId,
                               entropy_processor_factory)
from quadtree.common import EncodingInfo, get_dimension_bit_widths

BYTE_ORDER = "!"
PRESENT, MISSING = 1, 0


def header_struct(magic: bytes) -> struct.Struct:
    return struct.Struct(f"{BYTE_ORDER}{len(magic)}sfBBHHI")


class open_target:
    """`with open_target(path_or_file, mode) as f`: opens (and later closes) paths, passes file objects through."""

    def __init__(self, target, mode: str):
        self.target, self.mode, self.owned = target, mode, None

    def __enter__(self) -> BinaryIO:
        if isinstance(self.target, str):
            self.owned = open(self.target, self.mode)
            return self.owned
        return self.target

    def __exit__(self, *exc):
        if self.owned is not None:
            try:
                self.owned.close()
            except Exception:
                pass
        return False


def write_preamble(f: BinaryIO, bits: BitBuffer, magic: bytes, info: EncodingInfo, processor: EntropyProcessor,
                   coded: set[CodingDimension]) -> None:
    f.write(header_struct(magic).pack(magic, info.max_scale, info.scale_bits, info.offset_bits, info.img_width,
                                      info.img_height, 0))
    bits.write(processor.processor_id().value, ENTROPY_PROCESSOR_ID_BITS)
    for dim in CodingDimension:
        bits.write(PRESENT if dim in coded else MISSING, 1)
    for dim in CodingDimension:
        if dim in coded:
            processor.serialize_dimension_coding_structures(bits, dim)


def patch_size(f: BinaryIO, start: int, magic: bytes) -> None:
    end = f.tell()
    f.seek(start + header_struct(magic).size - 4)
    f.write(struct.pack(f"{BYTE_ORDER}I", end - start))
    f.seek(end)


def read_preamble(f: BinaryIO, bits: BitBuffer, magic: bytes, what: str):
    """Returns (EncodingInfo, total size, entropy processor ready for decoding)."""
    hs = header_struct(magic)
    raw = f.read(hs.size)
    if len(raw) != hs.size:
        raise Exception('Invalid file format: header is too short, empty file?')
    got_magic, max_scale, scale_bits, offset_bits, width, height, size = hs.unpack(raw)
    if got_magic != magic:
        raise Exception(f'Invalid file format: file is missing {what} magic')
    if scale_bits > BitBuffer.MAX_BIT_WIDTH or offset_bits > BitBuffer.MAX_BIT_WIDTH:
        # the header stores both as a byte; the reference's BitBuffer moves at most one 32-bit word per read
        raise Exception(f'Invalid file format: {scale_bits}/{offset_bits}-bit scale/offset fields exceed the '
                        f'{BitBuffer.MAX_BIT_WIDTH}-bit words of the bit buffer')
    info = EncodingInfo(scale_bits=scale_bits, offset_bits=offset_bits, max_scale=max_scale, img_width=width,
                        img_height=height)
    processor_id = bits.read(ENTROPY_PROCESSOR_ID_BITS)
    coded = {dim for dim in CodingDimension if bits.read(1) == PRESENT}
    processor = entropy_processor_factory(EntropyProcessorId(processor_id))   # ValueError on unknown ids, as the reference
    processor.prepare_for_decoding(get_dimension_bit_widths(info))
    for dim in CodingDimension:
        if dim in coded:
            processor.deserialize_dimension_coding_structures(bits, dim)
    return info, size, processor


def read_nodes_native(f: BinaryIO, bits: BitBuffer, processor: EntropyProcessor, fic_tree_kind: int, root_exps, cap: int):
    """The node section that follows the preamble, read by `fic_tree_read_nodes` (csrc/stream_host.cu) instead of one
    Python call per field: returns the preorder stream nodes[k,8] and leaves `f` where the reference's word-wise
    reader would have left it (the next stream of an RGB file starts there).  None if this entropy processor has no
    table form (the caller then runs the bit-by-bit loop)."""
    from entropy.processor import HuffmanEntropyProcessor
    if type(processor) is not HuffmanEntropyProcessor:
        return None
    window = bits.native_window()
    if window is None:
        return None
    import numpy as np

    import fic_b200
    start, data, bit_pos = window
    tables, rows = [], 0
    raw_width, huff_root = [], []
    for dim in CodingDimension:
        decoder = processor.dimension_decoders[dim.value]
        raw_width.append(int(processor.dimension_default_bit_widths[dim.value]))
        if decoder is None:
            huff_root.append(-1)
            continue
        t = decoder.table.copy()
        inner = t[:, 0] >= 0
        t[inner, 0] += rows              # rows of all trees share one table: shift the child links
        t[inner, 1] += rows
        huff_root.append(rows)
        rows += t.shape[0]
        tables.append(t)
    table = np.concatenate(tables) if tables else np.zeros((0, 3), np.int32)
    try:
        nodes, end_bit = fic_b200.tree_read_nodes(fic_tree_kind, data, bit_pos, root_exps, raw_width, huff_root, table, cap)
    except ValueError as e:
        bits.resume_at(start, data, bit_pos)
        if "buffer returned EOF" in str(e):     # the reference's `assert node_type != EOF` (serialization.py:188)
            raise AssertionError("Invalid size encoding, buffer returned EOF") from None
        raise
    bits.resume_at(start, data, end_bit)
    return nodes
