"""Colour wrapper of the quadtree codec (reference: quadtree/rgb.py:14-93): the image is converted to Y/Cb/Cr, each
plane goes through its own `QuadtreeEncoder` / `QuadtreeDecoder` (three independent grayscale problems -- on the
GPU simply three more calls of the same kernels), and a file is the three `.fqt` streams back to back."""
from __future__ import annotations

from typing import BinaryIO

import numpy as np

from _stream import open_target
from quadtree.common import ColoredQuadtreeImage
from quadtree.decoder import QuadtreeDecoder
from quadtree.encoder import QuadtreeEncoder
from quadtree.serialization import QuadtreeDeserializer, QuadtreeSerializer
from utils import ChromaSubsampling, rgb_to_ycbcr, ycbcr_to_rgb


class RGBQuadtreeEncoder:
    def __init__(self, y_encoder: QuadtreeEncoder, cb_encoder: QuadtreeEncoder, cr_encoder: QuadtreeEncoder = None) -> None:
        self.y_encoder = y_encoder
        self.cb_encoder = cb_encoder
        self.cr_encoder = cr_encoder if cr_encoder is not None else cb_encoder

    def encode(self, rgb_img: np.ndarray, chroma_subsampling: ChromaSubsampling = "4:2:0") -> ColoredQuadtreeImage:
        y, cb, cr = rgb_to_ycbcr(rgb_img, chroma_subsampling)
        planes = [np.ascontiguousarray(p) for p in (y, cb, cr)]
        return ColoredQuadtreeImage(self.y_encoder.encode(planes[0]), self.cb_encoder.encode(planes[1]),
                                    self.cr_encoder.encode(planes[2]))


class RGBQuadtreeDecoder:
    def __init__(self, y_decoder: QuadtreeDecoder, cb_decoder: QuadtreeDecoder = None, cr_decoder: QuadtreeDecoder = None) -> None:
        self.y_decoder = y_decoder
        self.cb_decoder = cb_decoder if cb_decoder is not None else y_decoder
        self.cr_decoder = cr_decoder if cr_decoder is not None else self.cb_decoder

    def decode(self, colored_img: ColoredQuadtreeImage, dimension_posprocessor=None) -> np.ndarray:
        y = self.y_decoder.decode(colored_img.encoded_Y, dimension_posprocessor)
        cb = self.cb_decoder.decode(colored_img.encoded_Cb, dimension_posprocessor)
        cr = self.cr_decoder.decode(colored_img.encoded_Cr, dimension_posprocessor)
        if y.shape == cb.shape == cr.shape:
            subsampling = "4:4:4"
        elif cb.shape == cr.shape:
            subsampling = "4:2:0"
        else:
            raise Exception("unsupported chroma subsampling")
        return ycbcr_to_rgb((y, cb, cr), subsampling)


class RGBQuadtreeSerializer:
    def __init__(self, serializer: QuadtreeSerializer) -> None:
        self.serializer = serializer
        self.file: BinaryIO = None

    def serialize(self, img: ColoredQuadtreeImage, output: str | BinaryIO):
        with open_target(output, 'wb') as f:
            self.file = f
            for plane in (img.encoded_Y, img.encoded_Cb, img.encoded_Cr):
                self.serializer.serialize(plane, f)


class RGBQuadtreeDeserializer:
    def __init__(self, deserializer: QuadtreeDeserializer) -> None:
        self.deserializer = deserializer

    def deserialize(self, input: str | BinaryIO) -> ColoredQuadtreeImage:
        with open_target(input, 'rb') as f:
            self.file = f
            y = self.deserializer.deserialize(f)
            cb = self.deserializer.deserialize(f)
            cr = self.deserializer.deserialize(f)
            return ColoredQuadtreeImage(y, cb, cr)
