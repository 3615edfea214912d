"""Colour wrapper of the quadtree codec (reference: quadtree/rgb.py:14-93): the image is converted to Y/Cb/Cr, each
plane goes through its own `QuadtreeEncoder` / `QuadtreeDecoder` (three independent grayscale problems), and a file is
the three `.fqt` streams back to back.  Here the pixels cross PCIe once: the RGB image is uploaded, `fic_rgb_to_ycbcr`
writes the three planes in device memory, the encoders read them there; the decoder leaves its three planes on the
device, `fic_ycbcr_to_rgb` recombines them and only the RGB result comes back."""
from __future__ import annotations

from typing import BinaryIO

import numpy as np

import fic_b200
from _stream import open_target
from quadtree.common import ColoredQuadtreeImage
from quadtree.decoder import QuadtreeDecoder
from quadtree.encoder import QuadtreeEncoder
from quadtree.serialization import QuadtreeDeserializer, QuadtreeSerializer
from utils import ChromaSubsampling, ycbcr_to_rgb


def _device():
    import torch
    h = fic_b200.default_handle()
    return torch, h, torch.device("cuda", h.device)


def _chroma_shape(N: int, M: int, chroma_subsampling) -> tuple[int, int]:
    if chroma_subsampling == "4:4:4":
        return N, M
    if chroma_subsampling == "4:2:0":
        return N // 2 + N % 2, M // 2 + M % 2
    raise Exception("Unsupported chroma subsampling")


class RGBQuadtreeEncoder:
    def __init__(self, y_encoder: QuadtreeEncoder, cb_encoder: QuadtreeEncoder, cr_encoder: QuadtreeEncoder = None) -> None:
        self.y_encoder = y_encoder
        self.cb_encoder = cb_encoder
        self.cr_encoder = cr_encoder if cr_encoder is not None else cb_encoder

    def encode(self, rgb_img: np.ndarray, chroma_subsampling: ChromaSubsampling = "4:2:0") -> ColoredQuadtreeImage:
        torch, h, dev = _device()
        rgb = np.ascontiguousarray(rgb_img, np.uint8)
        N, M, _ = rgb.shape
        cs = _chroma_shape(N, M, chroma_subsampling)
        d_rgb = torch.from_numpy(rgb).to(dev)
        planes = [torch.empty(shape, dtype=torch.uint8, device=dev) for shape in ((N, M), cs, cs)]
        torch.cuda.current_stream(dev).synchronize()
        fic_b200.rgb_to_ycbcr_dev(h, d_rgb.data_ptr(), N, M, chroma_subsampling, *(p.data_ptr() for p in planes))
        encoded = [enc.encode_device(p.data_ptr(), *p.shape)
                   for enc, p in zip((self.y_encoder, self.cb_encoder, self.cr_encoder), planes)]
        return ColoredQuadtreeImage(*encoded)


class RGBQuadtreeDecoder:
    def __init__(self, y_decoder: QuadtreeDecoder, cb_decoder: QuadtreeDecoder = None, cr_decoder: QuadtreeDecoder = None) -> None:
        self.y_decoder = y_decoder
        self.cb_decoder = cb_decoder if cb_decoder is not None else y_decoder
        self.cr_decoder = cr_decoder if cr_decoder is not None else self.cb_decoder

    def decode(self, colored_img: ColoredQuadtreeImage, dimension_posprocessor=None) -> np.ndarray:
        imgs = (colored_img.encoded_Y, colored_img.encoded_Cb, colored_img.encoded_Cr)
        decoders = (self.y_decoder, self.cb_decoder, self.cr_decoder)
        shapes = [(q.info.img_height, q.info.img_width) for q in imgs]
        if shapes[0] == shapes[1] == shapes[2]:
            subsampling = "4:4:4"
        elif shapes[1] == shapes[2]:
            subsampling = "4:2:0"
        else:
            raise Exception("unsupported chroma subsampling")
        on_device = dimension_posprocessor is None or \
            getattr(dimension_posprocessor, "has_default_settings", lambda: False)()
        if not on_device or any(d.on_domain_decoded is not None for d in decoders):
            # a caller-supplied filter works on host arrays (decoder.py:58-59): decode plane by plane, recombine on the GPU
            planes = [d.decode(q, dimension_posprocessor) for d, q in zip(decoders, imgs)]
            return ycbcr_to_rgb(tuple(planes), subsampling)
        torch, h, dev = _device()
        N, M = shapes[0]
        if shapes[1] != _chroma_shape(N, M, subsampling):
            raise ValueError(f"chroma planes {shapes[1]} do not belong to a {N}x{M} image")
        planes = [torch.empty(shape, dtype=torch.uint8, device=dev) for shape in shapes]
        for d, q, p in zip(decoders, imgs, planes):
            d.decode_planes_device([q], [p], dimension_posprocessor)
        d_rgb = torch.empty((N, M, 3), dtype=torch.uint8, device=dev)
        fic_b200.ycbcr_to_rgb_dev(h, *(p.data_ptr() for p in planes), N, M, subsampling, d_rgb.data_ptr())
        h.synchronize()
        return d_rgb.cpu().numpy()


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
