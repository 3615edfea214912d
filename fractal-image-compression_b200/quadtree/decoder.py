"""`QuadtreeDecoder` (reference: quadtree/decoder.py:13-116): Jacobi iteration of the leaf transforms in
float64 on the GPU -- all leaves in parallel per sweep, int-truncating pool, `.T` then `rot90`
(the reference's decoder-side composition), AUTO stop on the relative Frobenius change."""
from __future__ import annotations

from typing import Callable, Optional

import numpy as np

import fic_b200
from quadtree.common import QuadtreeImage, forest_to_leaves

AUTO = -1


def leaf_visiting_order(leaves: np.ndarray, width: int, height: int) -> np.ndarray:
    """Indices of `leaves` (rows of `forest_to_leaves`) in the order quadtree/decoder.py visits them: the roots in
    `_de_partition` order (:63-76), every root breadth first with the children queued TL, TR, BL, BR (:78-116) -- inside
    a root that is by depth, then by the Morton code of the leaf's corner (row bit above column bit)."""
    from quadtree.common import iter_root_squares
    L = np.asarray(leaves)
    row, col, size = L[:, 0].astype(np.int64), L[:, 1].astype(np.int64), L[:, 2].astype(np.int64)
    key = np.zeros(len(L), np.int64)

    def spread(v):
        v = v & 0xffff
        v = (v | (v << 8)) & 0x00ff00ff
        v = (v | (v << 4)) & 0x0f0f0f0f
        v = (v | (v << 2)) & 0x33333333
        v = (v | (v << 1)) & 0x55555555
        return v

    for r, (ri, rj, e) in enumerate(iter_root_squares(width, height)):
        side = 1 << e
        m = (row >= ri) & (row < ri + side) & (col >= rj) & (col < rj + side)
        depth = e - np.log2(size[m]).astype(np.int64)
        morton = (spread(row[m] - ri) << 1) | spread(col[m] - rj)
        key[m] = (np.int64(r) << 40) | (depth << 34) | morton
    return np.argsort(key, kind="stable")


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
            return self._decode_leaf_by_leaf(quadtree_img, postprocessor)
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

    def _decode_leaf_by_leaf(self, quadtree_img: QuadtreeImage, postprocessor=None) -> np.ndarray:
        """`on_domain_decoded(next_img, counter)` after every single leaf (decoder.py:108-110).  The sweeps are Jacobi
        sweeps -- a leaf reads the previous sweep's image only -- so the image the reference shows its callback after
        leaf k of sweep s is sweep s for the leaves visited so far and sweep s - 1 everywhere else.  The GPU decodes
        whole sweeps; this slow path replays them block by block on the host in the reference's visiting order (roots
        in `_de_partition` order, breadth first inside a root).  Before the first sweep the reference's `next_img` is
        uninitialised memory (`np.empty_like`); here it is zeros."""
        info = quadtree_img.info
        H, W = info.img_height, info.img_width
        leaves, so = forest_to_leaves(quadtree_img, self.use_quantized_values)
        # how many sweeps the decode runs (the AUTO rule needs the decode itself to say)
        _, done = fic_b200.tree_decode(fic_b200.TREE_QUADTREE, leaves, so, H, W, self.iterations,
                                       self.stop_on_relative_error, self.min_iterations, self.max_iterations)
        self.iterations_run = done
        order = leaf_visiting_order(leaves, W, H)
        self.next_img = np.zeros((H, W), np.float64)
        for sweep in range(1, done + 1):
            cur, _ = fic_b200.tree_decode(fic_b200.TREE_QUADTREE, leaves, so, H, W, sweep, out_f64=True)
            for k in order:
                i, j, size = int(leaves[k, 0]), int(leaves[k, 1]), int(leaves[k, 2])
                self.next_img[i:i + size, j:j + size] = cur[i:i + size, j:j + size]
                self.decoded_counter += 1
                self.on_domain_decoded(self.next_img, self.decoded_counter)
        out = self.next_img.copy()
        if self.log_stop and self.iterations == AUTO and done < self.max_iterations:
            print(f'Auto decoding stopped after {done - 1} iterations')
        if postprocessor is not None:
            out = np.asarray(postprocessor.postprocess(out, quadtree_img))
        self.img = out.clip(0., 255.).astype(np.uint8)
        return self.img

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
