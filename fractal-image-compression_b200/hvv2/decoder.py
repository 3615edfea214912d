"""`HVDecoder` (reference: hvv2/decoder.py:11-89): Jacobi sweeps over the HV leaves on the GPU; pooling adds the
four pixels in float32 (`average_subsample_jit`), orientation 1 is an up-down flip, rotation 2 a half turn."""
from __future__ import annotations

from typing import Optional

import numpy as np

import fic_b200
from hvv2.common import HVImage, tree_to_leaves
from quadtree.decoder import AUTO


def _numpy_keeps_float32_times_float64_scalar() -> bool:
    """`float32 array * np.float64 scalar`: NumPy 1.x (value-based casting, the reference's pin is numpy==1.26.1)
    keeps float32, NumPy >= 2 (NEP 50) promotes to float64.  hvv2/decoder.py:79 does exactly this multiplication."""
    return (np.zeros(1, np.float32) * np.float64(1.0)).dtype == np.float32


class HVDecoder:
    def __init__(self, iterations: int = AUTO, stop_on_relative_error: float = 5e-3, min_iterations=2,
                 max_iterations=10, log_stop=False, numpy1_promotion: Optional[bool] = None) -> None:
        """`numpy1_promotion` (not in the reference): which NumPy the decode should agree with in the one place where
        the two differ (see `_numpy_keeps_float32_times_float64_scalar`).  None = the NumPy installed next to this
        package, True = the reference's pinned NumPy 1.26 (float32 affine also with quantised values), False = NumPy 2.
        The committed goldens were made under NumPy 2.3."""
        self.iterations = iterations
        self.stop_on_relative_error = stop_on_relative_error
        self.min_iterations = min_iterations
        self.max_iterations = max_iterations
        self.log_stop = log_stop
        self.img: Optional[np.ndarray] = None
        self.next_img: Optional[np.ndarray] = None
        self.use_quantized_values = True
        self.iterations_run = 0
        self.numpy1_promotion = _numpy_keeps_float32_times_float64_scalar() if numpy1_promotion is None \
            else bool(numpy1_promotion)

    def using_not_quantized_values(self) -> "HVDecoder":
        self.use_quantized_values = False
        return self

    def decode(self, hv_img: HVImage) -> np.ndarray:
        info = hv_img.info
        leaves, so = tree_to_leaves(hv_img, self.use_quantized_values)
        # with un-quantised values the reference multiplies a float32 array by a Python float, which stays float32
        # under every NumPy (hvv2/decoder.py:79); with quantised values scale is an np.float64 scalar: float64 under
        # NumPy >= 2, still float32 under the reference's pinned NumPy 1.26
        out, done = fic_b200.tree_decode(fic_b200.TREE_HV, leaves, so, info.img_height, info.img_width, self.iterations,
                                         self.stop_on_relative_error, self.min_iterations, self.max_iterations,
                                         f32_affine=(not self.use_quantized_values) or self.numpy1_promotion)
        self.iterations_run = done
        if self.log_stop and self.iterations == AUTO and done < self.max_iterations:
            print(f'Auto decoding stopped after {done - 1} iterations')
        self.img = out
        return out
