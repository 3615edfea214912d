"""GPU: the rows next to the quadtree codec through the C ABI -- the deblocking post-processor (fused into the decoder
and standalone) and the colour-plane fan-out -- against the reference's golden outputs (tests/golden/post_golden.npz) and
the CPU oracle.  Bytes and float64 values bit-exact."""
import numpy as np
import pytest

from conftest import synthetic_image
from test_oracle_golden import qt_params
from test_post_oracle import DEBLOCK_CASES, DEBLOCK_FLOAT_CASES, RGB_CASES, deblock_input, visit_ordered_leaves

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", DEBLOCK_CASES)
def test_deblocking_matches_reference(fic, orc, golden, images, name):
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.postprocessing import DeblockingQuadtreePostprocessor
    g, gp = golden["qt_golden"], golden["post_golden"]
    shape = images[name].shape
    q, leaves, so, perm = visit_ordered_leaves(orc, g, name, shape)
    # class API: the stock filter runs fused into the decoder
    assert np.array_equal(QuadtreeDecoder().decode(q, DeblockingQuadtreePostprocessor()), gp[f"{name}_decdb"])
    assert np.array_equal(QuadtreeDecoder(iterations=4).decode(q, DeblockingQuadtreePostprocessor()), gp[f"{name}_decdb4"])
    # standalone entry point on an arbitrary float image; leaf order must not matter
    x = deblock_input(name, shape)
    want = gp[f"{name}_deblock_out"] if name in DEBLOCK_FLOAT_CASES else orc.qt_deblock(x, leaves[perm][:, :3])
    assert np.array_equal(fic.qt_deblock(x, leaves), want)
    assert np.array_equal(fic.qt_deblock(x, leaves[perm]), want)
    assert np.array_equal(fic.qt_deblock(x, leaves[::-1].copy()), want)
    assert np.array_equal(DeblockingQuadtreePostprocessor().postprocess(x, q), want)


def test_deblocking_custom_weights_and_user_postprocessor(fic, orc, golden, images):
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.postprocessing import DeblockingQuadtreePostprocessor, QuadtreePostprocessor
    name = "cauliflower240x92"
    g = golden["qt_golden"]
    shape = images[name].shape
    q, leaves, so, perm = visit_ordered_leaves(orc, g, name, shape)
    x = deblock_input(name, shape)
    post = DeblockingQuadtreePostprocessor()
    post.own_weight_max_depth, post.neigh_weight_max_depth = 0.7, 0.3
    post.own_weight_not_max_depth, post.neigh_weight_not_max_depth = 0.55, 0.45
    post.max_depth_range_size = 8
    assert not post.has_default_settings()
    want = orc.qt_deblock(x, leaves[perm][:, :3], (0.7, 0.3, 0.55, 0.45), 8)
    assert np.array_equal(post.postprocess(x, q), want)
    # non-default settings: the decoder hands the unclipped float image to the filter, as decoder.py:58-61 does
    flt, _ = fic.tree_decode(fic.TREE_QUADTREE, leaves, so, shape[0], shape[1], out_f64=True)
    assert flt.dtype == np.float64
    want8 = orc.qt_deblock(flt, leaves[perm][:, :3], (0.7, 0.3, 0.55, 0.45), 8).clip(0., 255.).astype(np.uint8)
    assert np.array_equal(QuadtreeDecoder().decode(q, post), want8)
    assert np.array_equal(flt.clip(0., 255.).astype(np.uint8), golden["qt_golden"][f"{name}_dec"])

    class Invert(QuadtreePostprocessor):
        def postprocess(self, decoded_img, quadtree_img):
            assert decoded_img.dtype == np.float64
            return 255.0 - decoded_img

    assert np.array_equal(QuadtreeDecoder().decode(q, Invert()), (255.0 - flt).clip(0., 255.).astype(np.uint8))


@pytest.mark.parametrize("seed,shape", [(1, (64, 64)), (2, (70, 45)), (3, (33, 96)), (4, (128, 40)), (5, (17, 23)), (6, (250, 131))])
def test_deblocking_and_decode_on_random_forests(fic, orc, seed, shape):
    """Random forests over ragged images with uneven depths: the visiting-order keys of the deblocking kernel (root index,
    depth, Morton code) against the oracle's sequential replay, standalone and fused into the decoder."""
    from quadtree.common import EncodingInfo, QuadtreeImage, forest_to_leaves
    from test_serialization import _random_quadtree_stream
    rng = np.random.default_rng(100 + seed)
    H, W = shape
    nodes, vals, n_roots = _random_quadtree_stream(rng, W, H, min_side=2, p_split=0.65)
    q = QuadtreeImage(EncodingInfo(7, 5, 1.0, W, H), stream=(nodes, vals, n_roots))
    leaves, so = forest_to_leaves(q)
    so[:, 0] *= 0.5                                           # keep the iteration tame
    rects, index = orc.qt_visit_order(nodes, n_roots, H, W)
    perm = (np.cumsum(nodes[:, 0] == 1) - 1)[index]
    assert np.array_equal(leaves[perm][:, :3], rects)
    x = rng.uniform(-30.0, 290.0, (H, W))
    want = orc.qt_deblock(x, rects)
    assert np.array_equal(fic.qt_deblock(x, leaves), want)
    assert np.array_equal(fic.qt_deblock(x, leaves[rng.permutation(len(leaves))]), want)
    for it in (3, fic.ITER_AUTO):
        got, n_it = fic.tree_decode(fic.TREE_QUADTREE, leaves, so, H, W, iterations=it, deblock=True)
        ref, n_ref = orc.tree_decode(0, leaves[perm], so[perm], H, W, iterations=it, deblock=True)
        assert n_it == n_ref and np.array_equal(got, ref)
        got, _ = fic.tree_decode(fic.TREE_QUADTREE, leaves, so, H, W, iterations=it)
        ref, _ = orc.tree_decode(0, leaves, so, H, W, iterations=it)
        assert np.array_equal(got, ref)


def test_deblocking_rejected_for_hv(fic):
    leaves = np.array([[0, 0, 4, 4, 0, 0, 0, 0]], np.int32)
    with pytest.raises(ValueError):
        fic.tree_decode(fic.TREE_HV, leaves, np.zeros((1, 2)), 8, 8, deblock=True)


@pytest.mark.parametrize("k", RGB_CASES)
def test_colour_conversion_matches_reference(fic, golden, k):
    import utils
    gp = golden["post_golden"]
    rgb = gp[f"rgb{k}"]
    N, M, _ = rgb.shape
    y, cb, cr = utils.rgb_to_ycbcr(rgb, "4:4:4")
    assert all(np.array_equal(a, gp[f"rgb{k}_444_{t}"]) for a, t in ((y, "y"), (cb, "cb"), (cr, "cr")))
    assert np.array_equal(utils.ycbcr_to_rgb((y, cb, cr), "4:4:4"), gp[f"rgb{k}_444_back"])
    if N % 2:
        with pytest.raises(ValueError):
            utils.rgb_to_ycbcr(rgb, "4:2:0")
        return
    y, cb, cr = utils.rgb_to_ycbcr(rgb, "4:2:0")
    assert all(np.array_equal(a, gp[f"rgb{k}_420_{t}"]) for a, t in ((y, "y"), (cb, "cb"), (cr, "cr")))
    if f"rgb{k}_420_back_raises" in gp.files:
        with pytest.raises(ValueError):
            utils.ycbcr_to_rgb((y, cb, cr), "4:2:0")
    else:
        assert np.array_equal(utils.ycbcr_to_rgb((y, cb, cr), "4:2:0"), gp[f"rgb{k}_420_back"])
    with pytest.raises(Exception):
        utils.rgb_to_ycbcr(rgb, "4:2:2")


@pytest.mark.parametrize("shape", [(1024, 1024), (516, 771)])
def test_colour_conversion_large_against_oracle(fic, orc, shape):
    import utils
    rgb = np.stack([synthetic_image(40 + c, *shape) for c in range(3)], axis=-1)
    for mode, sub in (("4:4:4", False), ("4:2:0", True)):
        got = utils.rgb_to_ycbcr(rgb, mode)
        want = orc.rgb_to_ycbcr(rgb, sub)
        assert all(np.array_equal(a, b) for a, b in zip(got, want))
        assert np.array_equal(utils.ycbcr_to_rgb(got, mode), orc.ycbcr_to_rgb(*want, sub))


@pytest.mark.parametrize("subsampling", ["4:4:4", "4:2:0"])
def test_rgb_quadtree_device_path_with_deblocking(fic, orc, subsampling):
    """RGBQuadtreeEncoder/Decoder keep the planes on the device; the result must equal the plane-by-plane host route
    (oracle colour conversion -> per-plane encode -> per-plane decode with the filter -> oracle recombination)."""
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.encoder import QuadtreeEncoder
    from quadtree.postprocessing import DeblockingQuadtreePostprocessor
    from quadtree.rgb import RGBQuadtreeDecoder, RGBQuadtreeEncoder
    rgb = np.stack([synthetic_image(s, 64, 96) for s in (21, 22, 23)], axis=-1)
    params = (7, 5, 1.0, False, False, 4, 16, 1.0)
    col = RGBQuadtreeEncoder(QuadtreeEncoder(*params), QuadtreeEncoder(*params)).encode(rgb, subsampling)
    planes = orc.rgb_to_ycbcr(rgb, subsampling == "4:2:0")
    decoded, decoded_db = [], []
    for plane, enc in zip(planes, (col.encoded_Y, col.encoded_Cb, col.encoded_Cr)):
        alone = QuadtreeEncoder(*params).encode(plane)
        assert (enc.info.img_height, enc.info.img_width) == plane.shape
        decoded.append(QuadtreeDecoder().decode(enc))
        decoded_db.append(QuadtreeDecoder().decode(enc, DeblockingQuadtreePostprocessor()))
        assert np.array_equal(decoded[-1], QuadtreeDecoder().decode(alone))
    dec = RGBQuadtreeDecoder(QuadtreeDecoder())
    assert np.array_equal(dec.decode(col), orc.ycbcr_to_rgb(*decoded, subsampling == "4:2:0"))
    assert np.array_equal(dec.decode(col, DeblockingQuadtreePostprocessor()),
                          orc.ycbcr_to_rgb(*decoded_db, subsampling == "4:2:0"))
    psnr = 10 * np.log10(255.0 ** 2 / np.mean((dec.decode(col).astype(float) - rgb) ** 2))
    assert psnr > 20.0
