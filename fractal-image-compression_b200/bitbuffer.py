"""MSB-first bit stream over big-endian 32-bit words -- the container of the reference's `.fqt` / HV files
(reference: bitbuffer.py:8-67).  Same observable behaviour as the reference class: a word is emitted the moment
32 bits have accumulated, `read` returns EOF (-1) when fewer than four bytes are left, `flush` zero-pads.
"""
from __future__ import annotations

import struct
from typing import BinaryIO, Optional

import numpy as np

EOF = -1


class BitBuffer:
    MAX_BIT_WIDTH = 32
    MAX_BYTE_WIDTH = MAX_BIT_WIDTH // 8

    def __init__(self, backing_buff: BinaryIO, struct_byte_order: str = "!") -> None:
        self.buff = 0                  # pending bits, right-aligned
        self.accumulated_size = 0      # how many of them
        self.backing_buff = backing_buff
        self.struct_byte_order = struct_byte_order
        self._word = struct.Struct(f"{struct_byte_order}I")

    # ---- writing ---------------------------------------------------------------------------------
    def write(self, v: int, bit_width: int) -> Optional[bytes]:
        """Append the low `bit_width` bits of `v`.  Returns the 4 bytes written to the backing buffer when this
        call completed a word, else None (bitbuffer.py:18-37)."""
        room = self.MAX_BIT_WIDTH - self.accumulated_size
        if bit_width < room:
            self.buff = (self.buff << bit_width) | (v & ((1 << bit_width) - 1))
            self.accumulated_size += bit_width
            return None
        spill = bit_width - room                       # bits of v that start the next word
        word = (self.buff << room) | ((v >> spill) & ((1 << room) - 1))
        packed = self._word.pack(word)
        self.buff = v & ((1 << spill) - 1)
        self.accumulated_size = spill
        self.backing_buff.write(packed)
        return packed

    def write_many(self, values, widths) -> None:
        """Append many fields at once: the same stream and the same final state as `write(v, w)` for every pair in
        order, packed by `fic_pack_fields` (one native pass instead of ~10^5 Python calls for a 512x512 tree)."""
        widths = np.asarray(widths, np.int64)
        values = np.asarray(values, np.uint64)
        if widths.size == 0:
            return
        if int(widths.max()) > 64 or int(widths.min()) < 0:
            raise ValueError("field widths must lie in 0..64")
        if self._word.pack(1) != b"\x00\x00\x00\x01":
            for v, w in zip(values.tolist(), widths.tolist()):     # little-endian containers: not a format the reference writes
                self.write(v, w)
            return
        import fic_b200
        data, self.buff, self.accumulated_size = fic_b200.pack_fields(values, widths, self.buff, self.accumulated_size)
        if data:
            self.backing_buff.write(data)

    def flush(self) -> bytes:
        """Zero-pad the pending bits to a full word and write it -- also when nothing is pending
        (bitbuffer.py:58-64; callers guard with is_empty())."""
        packed = self._word.pack(self.buff << (self.MAX_BIT_WIDTH - self.accumulated_size))
        self.buff = 0
        self.accumulated_size = 0
        self.backing_buff.write(packed)
        return packed

    # ---- reading ---------------------------------------------------------------------------------
    def read(self, bit_width: int) -> int:
        """Next `bit_width` bits as an unsigned integer, or EOF when the backing buffer runs dry
        (bitbuffer.py:39-56)."""
        out = 0
        need = bit_width
        if self.accumulated_size < bit_width:
            need = bit_width - self.accumulated_size
            out = (self.buff & ((1 << self.accumulated_size) - 1)) << need
            raw = self.backing_buff.read(self.MAX_BYTE_WIDTH)
            if len(raw) < self.MAX_BYTE_WIDTH:
                return EOF
            self.buff = self._word.unpack(raw)[0]
            self.accumulated_size = self.MAX_BIT_WIDTH
        self.accumulated_size -= need
        return out | ((self.buff >> self.accumulated_size) & ((1 << need) - 1))

    def is_empty(self) -> bool:
        return self.accumulated_size == 0

    # ---- hand-over to the native readers ---------------------------------------------------------
    def native_window(self):
        """(start, data, bit_pos) for the native stream readers, or None when the backing buffer cannot seek: `data` is
        the rest of the file from the word the next unread bit lies in, `bit_pos` that bit's offset inside `data`."""
        f = self.backing_buff
        try:
            if not f.seekable():
                return None
        except AttributeError:
            return None
        pending = self.accumulated_size
        start = f.tell() - (self.MAX_BYTE_WIDTH if pending else 0)
        f.seek(start)
        data = f.read()
        return start, data, (self.MAX_BIT_WIDTH - pending) if pending else 0

    def resume_at(self, start: int, data: bytes, end_bit: int) -> None:
        """Put buffer and file where bit-by-bit reading up to `end_bit` of a `native_window` would have left them."""
        words, rem = divmod(end_bit, self.MAX_BIT_WIDTH)
        if rem:
            at = words * self.MAX_BYTE_WIDTH
            self.buff = self._word.unpack(data[at:at + self.MAX_BYTE_WIDTH])[0]
            self.accumulated_size = self.MAX_BIT_WIDTH - rem
            words += 1
        else:
            self.buff, self.accumulated_size = 0, 0
        self.backing_buff.seek(start + words * self.MAX_BYTE_WIDTH)
