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


# ---- colour conversion (reference: utils.py:36-97) -- host side, not part of the search arithmetic -----------------
from typing import Literal

ChromaSubsampling = Literal["4:4:4", "4:2:0"]

_FWD = np.array([[65.738, 129.057, 25.064], [-37.945, -74.494, 112.439], [112.439, -94.154, -18.285]])
_FWD_OFF = np.array([16., 128., 128.])


def _pool2_cpu(plane: np.ndarray) -> np.ndarray:
    """2x2 mean of an even-sized integer plane, float64 (the value `average_subsample` returns, utils.py:13-25)."""
    p = plane.astype(np.int32)
    return (p[0::2, 0::2] + p[1::2, 0::2] + p[0::2, 1::2] + p[1::2, 1::2]) / 4


def rgb_to_ycbcr(rgb_img: np.ndarray, chroma_subsampling: ChromaSubsampling):
    """BT.601-style studio-swing conversion with the reference's /256 scaling and truncating uint8 cast; "4:2:0"
    halves both chroma planes (odd sizes keep the reference's edge handling, utils.py:48-63)."""
    N, M, _ = rgb_img.shape
    x = rgb_img.astype(np.float64) / 256.
    R, G, B = x[..., 0], x[..., 1], x[..., 2]
    Y = 16. + 65.738 * R + 129.057 * G + 25.064 * B
    Cb = 128. - 37.945 * R - 74.494 * G + 112.439 * B
    Cr = 128. + 112.439 * R - 94.154 * G - 18.285 * B
    ycc = np.stack((Y, Cb, Cr), axis=-1).clip(0., 255.).astype(np.uint8)
    if chroma_subsampling == "4:4:4":
        return ycc[..., 0], ycc[..., 1], ycc[..., 2]
    if chroma_subsampling != "4:2:0":
        raise Exception("Unsupported chroma subsampling")
    Ne, Me = N - N % 2, M - M % 2
    out = []
    for c in (1, 2):
        band = np.empty((N // 2 + N % 2, M // 2 + M % 2), dtype=np.uint8)
        band[:band.shape[0] - N % 2, :band.shape[1] - M % 2] = _pool2_cpu(ycc[:Ne, :Me, c])
        if M % 2 == 1:
            band[:band.shape[0] - N % 2, -1] = ((ycc[:Ne:2, -1, c].astype(np.int32) + ycc[1:Ne:2, -1, c].astype(np.int32)) / 2).astype(np.uint8)
        if N % 2 == 1:
            # utils.py:58 indexes `band[-1 :band.shape[1]-M%2]` (a row slice, not a row) -- kept verbatim
            band[-1:band.shape[1] - M % 2] = ((ycc[-1, :Me:2, c].astype(np.int32) + ycc[-1, 1:Me:2, c].astype(np.int32)) / 2).astype(np.uint8)
        if M % 2 == 1 and N % 2 == 1:
            band[-1, -1] = ycc[-1, -1, c]
        out.append(band)
    return ycc[..., 0], out[0], out[1]


def ycbcr_to_rgb(ycbcr_img, chroma_subsampling: ChromaSubsampling) -> np.ndarray:
    if chroma_subsampling == "4:4:4":
        planes = ycbcr_img
    elif chroma_subsampling == "4:2:0":
        N, M = ycbcr_img[0].shape
        Ns, Ms = ycbcr_img[1].shape
        Ne, Me = N - N % 2, M - M % 2
        ups = []
        for c in (1, 2):
            src = ycbcr_img[c]
            band = np.empty_like(ycbcr_img[0])
            core = src[:Ns - Ns % 2, :Ms - M % 2]
            band[:Ne:2, :Me:2] = core
            band[1:Ne:2, :Me:2] = core
            band[:Ne:2, 1:Me:2] = core
            band[1:Ne:2, 1:Me:2] = core
            if M % 2 == 1:
                band[:Ne:2, -1] = src[:, -1]
                band[1:Ne:2, -1] = src[:, -1]
            if N % 2 == 1:
                band[-1, :Me:2] = src[-1, :]
                band[-1, 1:Me:2] = src[:, -1]      # utils.py:84 (sic)
            if M % 2 == 1 and N % 2 == 1:
                band[-1, -1] = src[-1, -1]
            ups.append(band)
        planes = (ycbcr_img[0], ups[0], ups[1])
    else:
        raise Exception("Unsupported chroma subsampling")
    Y, Cb, Cr = np.array(planes, dtype=np.float64) / 256.
    R = 298.082 * Y + 408.583 * Cr - 222.921
    G = 298.082 * Y - 100.291 * Cb - 208.120 * Cr + 135.576
    B = 298.082 * Y + 516.412 * Cb - 276.836
    return np.stack((R, G, B), axis=-1).clip(0., 255.).astype(np.uint8)
