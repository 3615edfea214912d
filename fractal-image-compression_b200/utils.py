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
