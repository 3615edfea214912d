"""CPU: the oracle's restatement of the rows next to the quadtree codec -- DeblockingQuadtreePostprocessor
(quadtree/postprocessing.py) and the colour conversion (utils.py:36-97) -- against outputs of the UNMODIFIED reference
(tests/golden/post_golden.npz, made by tests/golden/make_golden_post.py).  Bit-exact."""
import numpy as np
import pytest

from test_oracle_golden import qt_params


def deblock_input(name, shape):
    """The seeded float64 image tests/golden/make_golden_post.py fed to the reference's post-processor."""
    return np.random.default_rng(len(name)).uniform(-20.0, 280.0, size=shape)

DEBLOCK_FLOAT_CASES = ["cauliflower64", "cauliflower128", "cauliflower240x92"]
DEBLOCK_CASES = ["cauliflower64", "cauliflower128", "cauliflower240x92", "cauliflower256", "lena313x413"]
RGB_CASES = list(range(8))


def visit_ordered_leaves(orc, g, name, img_shape):
    """(leaves, so) of a golden quadtree case in the order the reference visits the leaves."""
    from quadtree.common import EncodingInfo, QuadtreeImage, forest_to_leaves, stream_to_forest
    p = qt_params(g, name)
    nodes, vals, nroots = g[f"{name}_nodes"], g[f"{name}_vals"], int(g[f"{name}_nroots"])
    forest = stream_to_forest(np.c_[nodes, np.zeros(len(nodes), np.int32)], np.c_[vals, np.zeros(len(vals))], nroots)
    q = QuadtreeImage(EncodingInfo(p[0], p[1], p[2], img_shape[1], img_shape[0]), forest)
    leaves, so = forest_to_leaves(q)
    rects, index = orc.qt_visit_order(nodes, nroots, img_shape[0], img_shape[1])
    leaf_rank = np.cumsum(nodes[:, 0] == 1) - 1
    perm = leaf_rank[index]
    assert np.array_equal(leaves[perm][:, :3], rects)
    return q, leaves, so, perm


@pytest.mark.parametrize("name", DEBLOCK_CASES)
def test_deblocking_postprocessor(orc, golden, images, name):
    g, gp = golden["qt_golden"], golden["post_golden"]
    shape = images[name].shape
    _, leaves, so, perm = visit_ordered_leaves(orc, g, name, shape)
    if name in DEBLOCK_FLOAT_CASES:
        x = deblock_input(name, shape)
        assert np.array_equal(orc.qt_deblock(x, leaves[perm][:, :3]), gp[f"{name}_deblock_out"])
        # the result does depend on the visiting order
        assert not np.array_equal(orc.qt_deblock(x, leaves[:, :3]), gp[f"{name}_deblock_out"])
    out, _ = orc.tree_decode(0, leaves[perm], so[perm], shape[0], shape[1], deblock=True)
    assert np.array_equal(out, gp[f"{name}_decdb"])
    out, _ = orc.tree_decode(0, leaves[perm], so[perm], shape[0], shape[1], iterations=4, deblock=True)
    assert np.array_equal(out, gp[f"{name}_decdb4"])
    assert not np.array_equal(gp[f"{name}_decdb"], g[f"{name}_dec"])   # the filter is not a no-op on these images


@pytest.mark.parametrize("k", RGB_CASES)
def test_colour_conversion(orc, golden, k):
    gp = golden["post_golden"]
    rgb = gp[f"rgb{k}"]
    N, M, _ = rgb.shape
    y, cb, cr = orc.rgb_to_ycbcr(rgb, False)
    assert all(np.array_equal(a, gp[f"rgb{k}_444_{t}"]) for a, t in ((y, "y"), (cb, "cb"), (cr, "cr")))
    assert np.array_equal(orc.ycbcr_to_rgb(y, cb, cr, False), gp[f"rgb{k}_444_back"])
    if N % 2:
        # utils.py:58 writes a row slice: the reference raises or leaves the last chroma row uninitialised
        with pytest.raises(ValueError):
            orc.rgb_to_ycbcr(rgb, True)
        return
    y, cb, cr = orc.rgb_to_ycbcr(rgb, True)
    assert all(np.array_equal(a, gp[f"rgb{k}_420_{t}"]) for a, t in ((y, "y"), (cb, "cb"), (cr, "cr")))
    if f"rgb{k}_420_back_raises" in gp.files:
        assert N % 4 != 0
        with pytest.raises(ValueError):
            orc.ycbcr_to_rgb(y, cb, cr, True)
    else:
        assert np.array_equal(orc.ycbcr_to_rgb(y, cb, cr, True), gp[f"rgb{k}_420_back"])
