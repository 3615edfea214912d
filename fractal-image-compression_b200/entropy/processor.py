"""Per-dimension entropy coding of the serialised transform fields (reference: entropy/processor.py).

A coding dimension (domain row, domain column, quantised scale, quantised offset) is Huffman-coded only when that
saves at least one payload bit; the table cost is not counted (processor.py:102-106).  The processor is stateful
in the same way as the reference: encoders chosen in an earlier `prepare_for_encoding` call stay installed for
dimensions that are not worth coding in a later call (processor.py:88-98) -- relevant when one serialiser object is
reused for the Y, Cb and Cr planes (quadtree/rgb.py:62-64).
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from enum import Enum

from bitbuffer import BitBuffer
from entropy.abccodec import Code, CodeWidth, EntropyEncoder, IdentityEncoder
from entropy.huffman import (CodingTable, HuffmanDecoder, HuffmanDeserializer, HuffmanEncoder, HuffmanSerializer,
                             ordered_counts)


class CodingDimension(Enum):
    # values are list indices: 0..n-1 (processor.py:11-16)
    DOMAIN_ROW = 0
    DOMAIN_COL = 1
    SCALE = 2
    OFFSET = 3


ENTROPY_PROCESSOR_ID_BITS = 8


class EntropyProcessorId(Enum):
    HUFFMAN = 1


class EntropyProcessor(ABC):
    @abstractmethod
    def prepare_for_encoding(self, dimensions: dict[CodingDimension, list[int]],
                             dimension_default_bit_widths: list[int]) -> dict[CodingDimension, int]:
        """Build the coders; returns {dimension: payload bits saved} for the dimensions worth coding."""

    @abstractmethod
    def serialize_dimension_coding_structures(self, write_buff: BitBuffer, dimension: CodingDimension):
        ...

    @abstractmethod
    def encode(self, value: int, dimension: CodingDimension) -> tuple[Code, CodeWidth]:
        ...

    def encode_column(self, values, dimension: CodingDimension):
        """(code words uint64, widths int64) for a numpy column of one dimension's symbols; processors with per-dimension
        encoders override this with the encoder's own column method."""
        import numpy as np
        uniq, inverse = np.unique(np.asarray(values), return_inverse=True)
        pairs = [self.encode(int(v), dimension) for v in uniq]
        return (np.array([p[0] for p in pairs], np.uint64)[inverse], np.array([p[1] for p in pairs], np.int64)[inverse])

    @abstractmethod
    def prepare_for_decoding(self, dimension_default_bit_widths: list[int]):
        ...

    @abstractmethod
    def deserialize_dimension_coding_structures(self, read_buff: BitBuffer, dimension: CodingDimension):
        ...

    @abstractmethod
    def decode(self, read_buff: BitBuffer, dimension: CodingDimension) -> int:
        ...

    @abstractmethod
    def processor_id(self) -> EntropyProcessorId:
        ...


class HuffmanEntropyProcessor(EntropyProcessor):
    def __init__(self) -> None:
        n = len(CodingDimension)
        self.dimension_encoders: list[EntropyEncoder | None] = [None] * n
        self.dimension_decoders: list[HuffmanDecoder | None] = [None] * n
        self.dimension_default_bit_widths: list[int] = [-1] * n

    def prepare_for_encoding(self, dimensions, dimension_default_bit_widths):
        assert len(dimension_default_bit_widths) == len(self.dimension_encoders), "Every dimension requires default bit width"
        self.dimension_default_bit_widths = dimension_default_bit_widths
        gains: dict[CodingDimension, int] = {}
        for dim, values in dimensions.items():
            coder = HuffmanEncoder()
            table = coder.prepare_for_encoding(values)
            saved = self._payload_bits_saved(table, values, dimension_default_bit_widths[dim.value])
            if saved > 0:
                gains[dim] = saved
                self.dimension_encoders[dim.value] = coder
        for k, width in enumerate(dimension_default_bit_widths):
            if self.dimension_encoders[k] is None:      # only never-set slots: earlier coders are kept (see module doc)
                self.dimension_encoders[k] = IdentityEncoder(width)
        return gains

    @staticmethod
    def _payload_bits_saved(table: CodingTable, values: list[int], raw_width: int) -> int:
        return sum(count * (raw_width - table[sym][1]) for sym, count in ordered_counts(values))

    def serialize_dimension_coding_structures(self, write_buff, dimension):
        coder = self.dimension_encoders[dimension.value]
        assert isinstance(coder, HuffmanEncoder)
        HuffmanSerializer().serialize(coder.root, self.dimension_default_bit_widths[dimension.value], write_buff)

    def encode(self, value, dimension):
        return self.dimension_encoders[dimension.value].encode(value)

    def encode_column(self, values, dimension):
        return self.dimension_encoders[dimension.value].encode_column(values)

    def prepare_for_decoding(self, dimension_default_bit_widths):
        self.dimension_default_bit_widths = dimension_default_bit_widths

    def deserialize_dimension_coding_structures(self, read_buff, dimension):
        val_bits = self.dimension_default_bit_widths[dimension.value]
        window = read_buff.native_window()
        if window is not None:      # the tree dump is read by fic_huffman_read_tree; node objects only on demand
            import fic_b200
            start, data, bit_pos = window
            try:
                table, end_bit = fic_b200.huffman_read_tree(data, bit_pos, val_bits)
            except ValueError as e:
                read_buff.resume_at(start, data, bit_pos)
                if "buffer returned EOF" not in str(e):
                    raise
                table = None            # truncated file: let the bit-by-bit reader fail the way the reference does
            if table is not None:
                read_buff.resume_at(start, data, end_bit)
                self.dimension_decoders[dimension.value] = HuffmanDecoder(table=table)
                return
        root = HuffmanDeserializer().deserialize(read_buff, val_bits)
        self.dimension_decoders[dimension.value] = HuffmanDecoder(root)

    def decode(self, read_buff, dimension):
        decoder = self.dimension_decoders[dimension.value]
        if decoder is not None:
            return decoder.decode(read_buff)
        return read_buff.read(self.dimension_default_bit_widths[dimension.value])

    def processor_id(self):
        return EntropyProcessorId.HUFFMAN


def entropy_processor_factory(id: EntropyProcessorId) -> EntropyProcessor:
    if id == EntropyProcessorId.HUFFMAN:
        return HuffmanEntropyProcessor()
    raise Exception(f"Unknown entropy processor: {id}")
