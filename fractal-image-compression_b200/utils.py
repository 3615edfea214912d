"""Image helpers of the reference's `utils` module (utils.py:9-34).  The two pooling functions run on the GPU
(`fic_average_subsample`); they exist for API parity -- the codecs pool inside their own kernels."""
from __future__ import annotations

import numpy as np

import fic_b200


def load_grayscale(src_path: str) -> np.ndarray:
    from PIL import Image
    with Image.open(src_path) as im:
        return np.array(im.convert("L"))


def average_subsample(arr: np.ndarray) -> np.ndarray:
    """2x2 box mean with the reference's int32 truncation of the input (utils.py:13-25); float64 result."""
    a = np.asarray(arr)
    return fic_b200.average_subsample(a.astype(np.float64), jit=False)


def average_subsample_jit(arr: np.ndarray) -> np.ndarray:
    """2x2 box mean with float32 adds (utils.py:27-34); float32 result."""
    a = np.asarray(arr)
    return fic_b200.average_subsample(a.astype(np.float64), jit=True).astype(np.float32)


# ---- colour conversion (reference: utils.py:36-97), on the device: csrc/colour.cu ---------------------------------
from typing import Literal

ChromaSubsampling = Literal["4:4:4", "4:2:0"]


def rgb_to_ycbcr(rgb_img: np.ndarray, chroma_subsampling: ChromaSubsampling):
    """uint8 [N,M,3] -> (Y, Cb, Cr) uint8 planes; "4:2:0" halves both chroma planes (2x2 integer mean, truncated; an odd
    last column keeps the reference's vertical-pair rule).  Raises ValueError for "4:2:0" with an odd height, where the
    reference (utils.py:57-61) raises or leaves the last chroma row uninitialised."""
    return fic_b200.rgb_to_ycbcr(rgb_img, chroma_subsampling)


def ycbcr_to_rgb(ycbcr_img, chroma_subsampling: ChromaSubsampling) -> np.ndarray:
    """(Y, Cb, Cr) -> uint8 [N,M,3]; "4:2:0" repeats every chroma sample over its 2x2 cell.  Raises ValueError where the
    reference does (utils.py:73-76: a height that is not a multiple of 4)."""
    y, cb, cr = ycbcr_img
    return fic_b200.ycbcr_to_rgb(y, cb, cr, chroma_subsampling)
