#!/usr/bin/env python
"""Golden fixtures for the rows either side of the quadtree codec, made by running the
UNMODIFIED reference (build container only: needs /root/reference, numba):

    python tests/golden/make_golden_post.py

post_golden.npz (committed):
  <case>_decdb / _decdb4   QuadtreeDecoder(AUTO / iterations=4).decode(q, DeblockingQuadtreePostprocessor()) of the
                           quadtree cases of qt_golden.npz (re-encoded here by the reference; the node streams are
                           checked against the stored ones first)
  <case>_deblock_out       DeblockingQuadtreePostprocessor().postprocess(deblock_input(case), q), small cases only
  rgb<k>                   seeded RGB inputs; rgb<k>_<444|420>_{y,cb,cr} = utils.rgb_to_ycbcr, rgb<k>_<444|420>_back =
                           utils.ycbcr_to_rgb of those planes; rgb<k>_<mode>_raises = 1 where the reference raises
"""
import os
import sys

import numpy as np

REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "src"))
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


def deblock_input(name, shape):
    """Seeded float64 image with values outside 0..255 too (the filter runs before the clip)."""
    return np.random.default_rng(len(name)).uniform(-20.0, 280.0, size=shape)


def main():
    from make_golden import qt_flatten
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.encoder import QuadtreeEncoder
    from quadtree.postprocessing import DeblockingQuadtreePostprocessor
    from utils import rgb_to_ycbcr, ycbcr_to_rgb

    imgs = np.load(os.path.join(HERE, "images.npz"))
    g = np.load(os.path.join(HERE, "qt_golden.npz"))
    out = {}
    for name in ("cauliflower64", "cauliflower128", "cauliflower240x92", "cauliflower256", "lena313x413"):
        p = g[f"{name}_params"]
        params = (int(p[0]), int(p[1]), float(p[2]), bool(p[3]), bool(p[4]), int(p[5]), int(p[6]), float(p[7]))
        q = QuadtreeEncoder(*params).encode(imgs[name])
        nodes, vals, nroots = qt_flatten(q)
        assert np.array_equal(nodes, g[f"{name}_nodes"]) and np.array_equal(vals, g[f"{name}_vals"]), name
        out[f"{name}_decdb"] = QuadtreeDecoder().decode(q, DeblockingQuadtreePostprocessor())
        out[f"{name}_decdb4"] = QuadtreeDecoder(iterations=4).decode(q, DeblockingQuadtreePostprocessor())
        if imgs[name].size <= 240 * 92:   # float64 fixtures only for the small cases (the input is re-derived from its seed)
            x = deblock_input(name, imgs[name].shape)
            out[f"{name}_deblock_out"] = DeblockingQuadtreePostprocessor().postprocess(x, q)
        print(name, "deblocked")

    shapes = [(16, 24), (64, 48), (12, 17), (8, 9), (6, 8), (9, 8), (7, 7), (40, 33)]
    for k, (N, M) in enumerate(shapes):
        rgb = np.random.default_rng(100 + k).integers(0, 256, size=(N, M, 3), dtype=np.uint8)
        if k == 0:
            rgb[:4] = 255
            rgb[4:8] = 0
        out[f"rgb{k}"] = rgb
        for mode, tag in (("4:4:4", "444"), ("4:2:0", "420")):
            try:
                with np.errstate(all="ignore"):
                    y, cb, cr = rgb_to_ycbcr(rgb, mode)
            except Exception as e:  # noqa: BLE001 - the reference's own failure modes are the fixture
                out[f"rgb{k}_{tag}_raises"] = np.array(1)
                print("rgb", k, (N, M), mode, "forward raises", type(e).__name__)
                continue
            out[f"rgb{k}_{tag}_y"], out[f"rgb{k}_{tag}_cb"], out[f"rgb{k}_{tag}_cr"] = y, cb, cr
            try:
                out[f"rgb{k}_{tag}_back"] = ycbcr_to_rgb((y, cb, cr), mode)
            except Exception as e:  # noqa: BLE001
                out[f"rgb{k}_{tag}_back_raises"] = np.array(1)
                print("rgb", k, (N, M), mode, "backward raises", type(e).__name__)
    np.savez_compressed(os.path.join(HERE, "post_golden.npz"), **out)
    print("done", len(out), "arrays")


if __name__ == "__main__":
    main()
