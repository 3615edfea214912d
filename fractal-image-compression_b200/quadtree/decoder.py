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
        # decoder.py:58-61: a post-processor sees the float image, the clip comes after it.  The stock deblocking
        # filter runs fused on the device; any other QuadtreePostprocessor is handed the float64 image as in the reference.
        fused = postprocessor is not None and getattr(postprocessor, "has_default_settings", lambda: False)()
        custom = postprocessor is not None and not fused
        out, done = fic_b200.tree_decode(fic_b200.TREE_QUADTREE, leaves, so, info.img_height, info.img_width,
                                         self.iterations, self.stop_on_relative_error, self.min_iterations,
                                         self.max_iterations, deblock=fused, out_f64=custom)
        self.iterations_run = done
        if self.log_stop and self.iterations == AUTO and done < self.max_iterations:
            print(f'Auto decoding stopped after {done - 1} iterations')
        if custom:
            out = np.asarray(postprocessor.postprocess(out, quadtree_img)).clip(0., 255.).astype(np.uint8)
        self.img = out
        return out

    def decode_planes_device(self, quadtree_imgs, out_tensors, postprocessor=None) -> None:
        """Decode several images into uint8 device tensors (torch, one per image) without a host round trip of the
        pixels -- the colour wrapper's path (quadtree/rgb.py:35-48).  Only the stock deblocking filter can run here."""
        import torch
        if postprocessor is not None and not getattr(postprocessor, "has_default_settings", lambda: False)():
            raise ValueError("only the stock DeblockingQuadtreePostprocessor runs on the device")
        h = fic_b200.default_handle(out_tensors[0].device.index)
        for qimg, out in zip(quadtree_imgs, out_tensors):
            info = qimg.info
            leaves, so = forest_to_leaves(qimg, self.use_quantized_values)
            d_leaves = torch.from_numpy(leaves).to(out.device)
            d_so = torch.from_numpy(so).to(out.device)
            torch.cuda.current_stream(out.device).synchronize()
            self.iterations_run = fic_b200.tree_decode_dev(
                h, fic_b200.TREE_QUADTREE, d_leaves.data_ptr(), d_so.data_ptr(), leaves.shape[0], info.img_height,
                info.img_width, out.data_ptr(), self.iterations, self.stop_on_relative_error, self.min_iterations,
                self.max_iterations, flags=fic_b200.QT_DEBLOCK if postprocessor is not None else 0)
            h.synchronize()
