"""`QuadtreeDecoder` (reference: quadtree/decoder.py:13-116): Jacobi iteration of the leaf transforms in
float64 on the GPU -- all leaves in parallel per sweep, int-truncating pool, `.T` then `rot90`
(the reference's decoder-side composition), AUTO stop on the relative Frobenius change."""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

import fic_b200
from quadtree.common import QuadtreeImage, forest_to_leaves

AUTO = -1


class QuadtreeDecoder:
    def __init__(self, iterations: int = AUTO, stop_on_relative_error: float = 5e-3, min_iterations=2,
                 max_iterations=10, log_stop=False,
                 on_domain_decoded: Optional[Callable[[np.ndarray, int], None]] = None) -> None:
        self.iterations = iterations
        self.stop_on_relative_error = stop_on_relative_error
        self.min_iterations = min_iterations
        self.max_iterations = max_iterations
        self.log_stop = log_stop
        self.img: Optional[np.ndarray] = None
        self.next_img: Optional[np.ndarray] = None
        self.use_quantized_values = True
        self.on_domain_decoded = on_domain_decoded
        self.decoded_counter = 0
        self.iterations_run = 0

    def using_not_quantized_values(self) -> "QuadtreeDecoder":
        self.use_quantized_values = False
        return self

    def decode(self, quadtree_img: QuadtreeImage, postprocessor=None) -> np.ndarray:
        if self.on_domain_decoded is not None:
            raise NotImplementedError(
                "on_domain_decoded observes the image after every single leaf (decoder.py:108-110); the GPU decoder "
                "updates all leaves of a sweep at once")
        info = quadtree_img.info
        leaves, so = forest_to_leaves(quadtree_img, self.use_quantized_values)
        out, done = fic_b200.tree_decode(fic_b200.TREE_QUADTREE, leaves, so, info.img_height, info.img_width,
                                         self.iterations, self.stop_on_relative_error, self.min_iterations,
                                         self.max_iterations)
        self.iterations_run = done
        if self.log_stop and self.iterations == AUTO and done < self.max_iterations:
            print(f'Auto decoding stopped after {done - 1} iterations')
        if postprocessor is not None:
            # the reference post-filters the float image before clipping (decoder.py:58-61)
            out = postprocessor.postprocess(out.astype(np.float64), quadtree_img).clip(0., 255.).astype(np.uint8)
        self.img = out
        return out
