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
