"""`Decoder`: iterates the stored transforms (bruteforce/decoder.py:7-32) on the GPU, including the
reference's quirks -- buffers alias after the first sweep, the pool truncates to int32, the final
uint8 cast does not clip, range 8 / domain 16 are hard-coded."""
from __future__ import annotations

import numpy as np

import fic_b200
from bruteforce.model import EncodedImage


class Decoder:
    def __init__(self) -> None:
        self.range_size = 8
        self.domain_size = 16

    def decode(self, encoded_img: EncodedImage, iterations=32) -> np.ndarray:
        idx, so = encoded_img.to_arrays()
        return fic_b200.bf_decode(idx, so, encoded_img.width, encoded_img.height, iterations)
