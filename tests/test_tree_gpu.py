"""GPU: quadtree and HV encoders/decoders through the C ABI against the reference's golden outputs
(tests/golden) and the CPU oracle.  All integer fields, float64 scale/offset and decoded bytes bit-exact."""
import numpy as np
import pytest

from conftest import synthetic_image
from test_oracle_golden import HV_CASES, QT_CASES, hv_params, qt_image, qt_params

pytestmark = pytest.mark.gpu


def qt_flat(qimg):
    nodes, vals = [], []
    for root in qimg.forest:
        stack = [root]
        while stack:
            n = stack.pop()
            if n.is_leaf():
                d = n.domain
                nodes.append([1, d.start_i, d.start_j, d.orientation, d.rotation, d.quantized_scale, d.quantized_offset])
                vals.append([d.scale, d.offset])
            else:
                nodes.append([0] * 7)
                vals.append([0.0, 0.0])
                stack.extend(reversed(n.children))
    return np.array(nodes, np.int32), np.array(vals, np.float64)


def hv_flat(root):
    nodes, vals = [], []
    stack = [root]
    while stack:
        n = stack.pop()
        if n.is_leaf():
            d = n.domain
            nodes.append([1, d.start_i, d.start_j, d.orientation, d.rotation, d.quantized_scale, d.quantized_offset])
            vals.append([d.scale, d.offset])
        else:
            nodes.append([0, int(n.split_info.vertical_split), n.split_info.split_idx, 0, 0, 0, 0])
            vals.append([0.0, 0.0])
            stack.append(n.c2)
            stack.append(n.c1)
    return np.array(nodes, np.int32), np.array(vals, np.float64)


@pytest.mark.parametrize("name", QT_CASES)
def test_quadtree_encoder_and_decoder_match_reference(fic, golden, images, name):
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.encoder import QuadtreeEncoder
    g = golden["qt_golden"]
    img = qt_image(images, name)
    q = QuadtreeEncoder(*qt_params(g, name)).encode(img)
    nodes, vals = qt_flat(q)
    assert len(q.forest) == int(g[f"{name}_nroots"])
    assert np.array_equal(nodes, g[f"{name}_nodes"])
    assert np.array_equal(vals, g[f"{name}_vals"])
    assert (q.info.img_width, q.info.img_height) == (img.shape[1], img.shape[0])
    assert np.array_equal(QuadtreeDecoder().decode(q), g[f"{name}_dec"])
    if f"{name}_dec4" in g.files:
        assert np.array_equal(QuadtreeDecoder(iterations=4).decode(q), g[f"{name}_dec4"])


@pytest.mark.parametrize("params", [(7, 5, 1.0, False, False, 4, 8, 2.0), (6, 6, 1.0, True, False, 2, 8, 1.0),
                                    (7, 5, 0.75, False, True, 2, 16, 0.5, 2, 4),
                                    (7, 5, 1.0, False, False, 4, 8, 2.0, 1, 1), (6, 6, 1.0, True, False, 2, 8, 1.0, 1, 1)],
                         ids=["default", "full_class_search", "any_domain_steps_2_4", "dense_pools", "dense_full_class"])
@pytest.mark.parametrize("seed,shape", [(1, (64, 64)), (2, (70, 45)), (3, (33, 96))])
def test_quadtree_against_oracle(fic, orc, params, seed, shape):
    img = synthetic_image(seed, *shape)
    want = orc.qt_encode(img, *params)
    got = fic.qt_encode(img, *params)
    assert got[2] == want[2]
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[1], want[1])


def test_rgb_quadtree_is_three_independent_planes(fic, orc):
    """the colour wrapper (reference quadtree/rgb.py:14-56) is three grayscale problems; every plane
    of the fan-out must equal the oracle's encode of that plane and the decode must be the planes' decodes recombined."""
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.encoder import QuadtreeEncoder
    from quadtree.rgb import RGBQuadtreeDecoder, RGBQuadtreeEncoder
    from utils import rgb_to_ycbcr, ycbcr_to_rgb
    rgb = np.stack([synthetic_image(s, 64, 96) for s in (11, 12, 13)], axis=-1)
    params = (7, 5, 1.0, False, False, 4, 16, 0.5)
    for subsampling in ("4:4:4", "4:2:0"):
        col = RGBQuadtreeEncoder(QuadtreeEncoder(*params), QuadtreeEncoder(*params)).encode(rgb, subsampling)
        planes = rgb_to_ycbcr(rgb, subsampling)
        decoded = []
        for plane, enc in zip(planes, (col.encoded_Y, col.encoded_Cb, col.encoded_Cr)):
            plane = np.ascontiguousarray(plane)
            want = orc.qt_encode(plane, *params)
            nodes, vals = qt_flat(enc)
            leaf = want[0][:, 0] == 1
            assert len(enc.forest) == want[2] and np.array_equal(nodes[:, 0] == 1, leaf)
            assert np.array_equal(nodes[leaf], want[0][leaf, :7]) and np.array_equal(vals[leaf], want[1][leaf, :2])
            decoded.append(QuadtreeDecoder().decode(enc))
            assert decoded[-1].shape == plane.shape
        out = RGBQuadtreeDecoder(QuadtreeDecoder()).decode(col)
        assert out.shape == rgb.shape and out.dtype == np.uint8
        assert np.array_equal(out, ycbcr_to_rgb(tuple(decoded), subsampling))


def test_quadtree_flat_image_and_placeholders(fic, orc):
    img = np.full((64, 64), 200, np.uint8)
    img[16:24, 40:48] = 13
    params = (7, 5, 1.0, False, False, 4, 16, 0.5)
    want = orc.qt_encode(img, *params)
    got = fic.qt_encode(img, *params)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_quadtree_domain_does_not_fit_raises(fic):
    from quadtree.encoder import QuadtreeEncoder
    with pytest.raises(ValueError):
        QuadtreeEncoder(7, 5, 1.0, False, False, 4, 64, 2.0).encode(synthetic_image(1, 64, 64))


def test_classify_against_oracle(fic, orc):
    img = synthetic_image(8, 96, 80)
    rng = np.random.default_rng(0)
    blocks = []
    for _ in range(300):
        side = int(2 ** rng.integers(1, 6))
        blocks.append((int(rng.integers(0, 96 - side + 1)), int(rng.integers(0, 80 - side + 1)), side, int(rng.integers(0, 2))))
    got = fic.qt_classify(img, blocks)
    for b, gt in zip(blocks, got):
        assert tuple(gt.tolist()) == orc.qt_classify(img, b[0], b[1], b[2], bool(b[3]))


@pytest.mark.parametrize("name", HV_CASES)
def test_hv_encoder_and_decoder_match_reference(fic, golden, name):
    from hvv2.decoder import HVDecoder
    from hvv2.encoder import HVEncoder
    g = golden["hv_golden"]
    img = g[f"{name}_img"]
    h = HVEncoder(*hv_params(g, name)).encode(img)
    nodes, vals = hv_flat(h.root)
    assert np.array_equal(nodes, g[f"{name}_nodes"])
    assert np.array_equal(vals, g[f"{name}_vals"])
    assert np.array_equal(HVDecoder().decode(h), g[f"{name}_dec"])
    assert np.array_equal(HVDecoder(iterations=5).using_not_quantized_values().decode(h), g[f"{name}_dec5nq"])


@pytest.mark.parametrize("name", ["lena256", "lena512"])
def test_hv_large_images_match_reference(fic, golden, name):
    from hvv2.decoder import HVDecoder
    from hvv2.encoder import HVEncoder
    g = golden["hv_golden"]
    if f"{name}_nodes" not in g.files:
        pytest.skip("slow HV golden not generated")
    h = HVEncoder(*hv_params(g, name)).encode(g[f"{name}_img"])
    nodes, vals = hv_flat(h.root)
    assert np.array_equal(nodes, g[f"{name}_nodes"])
    assert np.array_equal(vals, g[f"{name}_vals"])
    assert np.array_equal(HVDecoder().decode(h), g[f"{name}_dec"])


def test_hv_search_against_oracle(fic, orc):
    img = synthetic_image(6, 64, 80)
    rects = [(0, 0, 4, 4), (10, 7, 1, 4), (3, 60, 5, 1), (20, 20, 16, 16), (40, 8, 13, 7), (0, 0, 40, 32), (5, 5, 39, 2)]
    rec, vals = fic.hv_search(img, rects, 5, 7, 1.0)
    for (i, j, w, h), r, v in zip(rects, rec, vals):
        wr, wv = orc.hv_search(img, i, j, w, h, 5, 7, 1.0)
        assert np.array_equal(r, wr)
        assert np.array_equal(v, wv)


def test_hv_search_large_rectangles_and_chunked_candidates(fic, orc):
    """Rectangles above 4 096 pixels are read straight from the image instead of shared memory, and a call with few
    rectangles cuts each candidate list into chunks over several CTAs (hv_pick_kernel merges them): both against the
    oracle, including a rectangle with a single domain position and one with none."""
    img = synthetic_image(9, 288, 320)
    rects = [(0, 0, 70, 66), (16, 8, 96, 80), (100, 40, 159, 100), (3, 5, 9, 143), (0, 0, 160, 144), (7, 7, 3, 2)]
    rec, vals = fic.hv_search(img, rects, 5, 7, 1.0)
    for (i, j, w, h), r, v in zip(rects, rec, vals):
        wr, wv = orc.hv_search(img, i, j, w, h, 5, 7, 1.0)
        assert np.array_equal(r, wr), (i, j, w, h)
        assert np.array_equal(v, wv), (i, j, w, h)


@pytest.mark.parametrize("params", [(4, 4, 16, 16, 5, 7, 25.0), (1, 1, 8, 8, 6, 6, 4.0, 0.9), (2, 3, 64, 64, 5, 7, 60.0)])
def test_hv_against_oracle(fic, orc, params):
    img = synthetic_image(12, 56, 72)
    want = orc.hv_encode(img, *params)
    got = fic.hv_encode(img, *params)
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[1], want[1])


def test_tree_decoder_argument_checks(fic):
    leaves = np.array([[0, 0, 8, 8, 60, 60, 0, 0]], np.int32)
    with pytest.raises(ValueError):
        fic.tree_decode(fic.TREE_QUADTREE, leaves, np.zeros((1, 2)), 64, 64)


@pytest.mark.parametrize("seed,shape", [(1, (64, 64)), (2, (70, 45)), (3, (33, 96)), (4, (200, 120))])
def test_hv_decode_on_random_trees(fic, orc, seed, shape):
    """Random HV trees (arbitrary split positions, ragged images): leaf table from the flat stream (`fic_hv_layout`), Jacobi
    decode on the GPU against the oracle, quantised and float32-affine variants, fixed and AUTO sweeps."""
    from hvv2.common import HVImage, tree_to_leaves
    from quadtree.common import EncodingInfo
    from test_serialization import _random_hv_stream
    rng = np.random.default_rng(200 + seed)
    H, W = shape
    nodes, vals = _random_hv_stream(rng, W, H)
    vals[:, 0] = np.clip(vals[:, 0], -0.9, 0.9)
    vals[:, 1] = rng.uniform(0, 255, len(vals))
    hv = HVImage(EncodingInfo(5, 7, 1.0, W, H), stream=(nodes, vals))
    for quantised in (True, False):
        leaves, so = tree_to_leaves(hv, quantised)
        for it in (2, 6, fic.ITER_AUTO):
            got, n_it = fic.tree_decode(fic.TREE_HV, leaves, so, H, W, iterations=it, f32_affine=not quantised)
            ref, n_ref = orc.tree_decode(1, leaves, so, H, W, iterations=it, f32_affine=not quantised)
            assert n_it == n_ref and np.array_equal(got, ref)


def test_hv_decoder_follows_either_numpy_promotion_rule(fic, orc, golden):
    """hvv2/decoder.py:79 multiplies a float32 array by an np.float64 scalar: float64 under NumPy >= 2 (the goldens),
    float32 under the reference's pinned NumPy 1.26.  The class takes the choice explicitly; both against the oracle."""
    from hvv2.common import tree_to_leaves
    from hvv2.decoder import HVDecoder
    from hvv2.encoder import HVEncoder
    img = synthetic_image(21, 72, 88)
    hv = HVEncoder(4, 4, 16, 16, 5, 7, 25.0).encode(img)
    leaves, so = tree_to_leaves(hv, True)
    for np1 in (False, True):
        dec = HVDecoder(iterations=6, numpy1_promotion=np1)
        want, _ = orc.tree_decode(1, leaves, so, 72, 88, iterations=6, f32_affine=np1)
        assert np.array_equal(dec.decode(hv), want)
    auto = HVDecoder()          # default: the installed NumPy's rule
    assert auto.numpy1_promotion == ((np.zeros(1, np.float32) * np.float64(1.0)).dtype == np.float32)


def _reference_style_decode_with_callback(leaves_in_order, so_in_order, H, W, sweeps):
    """quadtree/decoder.py:34-116 restated with NumPy for a handful of leaves: Jacobi sweeps, `.T` then `rot90`, the
    int-truncating pool of utils.average_subsample, one snapshot of `next_img` after every leaf (zeros where the
    reference's `np.empty_like` buffer holds whatever memory it was given)."""
    img = np.zeros((H, W), np.float64)
    nxt = np.zeros((H, W), np.float64)
    shots = []
    for _ in range(sweeps):
        for (ri, rj, size, _, di, dj, orient, rot), (scale, offset) in zip(leaves_in_order.tolist(), so_in_order.tolist()):
            dom = np.copy(img[di:di + 2 * size, dj:dj + 2 * size])
            if orient == 1:
                dom = dom.T
            dom = np.rot90(dom, rot)
            d = dom.reshape(-1).reshape((-1, dom.shape[1] * 2)).astype(np.int32)
            x = d[:, :dom.shape[1]] + d[:, dom.shape[1]:]
            x = x[:, ::2] + x[:, 1::2]
            nxt[ri:ri + size, rj:rj + size] = x / 4 * scale + offset
            shots.append(nxt.copy())
        img[:] = nxt
    return shots


@pytest.mark.parametrize("shape", [(32, 32), (40, 52)])
def test_quadtree_decoder_leaf_by_leaf_callback(fic, orc, shape):
    """`on_domain_decoded` (decoder.py:16,27,108-110): the callback sees `next_img` after every single leaf, in the
    reference's visiting order; checked against a NumPy restatement of the reference's loop, ragged image included."""
    from quadtree.common import forest_to_leaves
    from quadtree.decoder import QuadtreeDecoder, leaf_visiting_order
    from quadtree.encoder import QuadtreeEncoder
    H, W = shape
    img = synthetic_image(33, H, W)
    q = QuadtreeEncoder(7, 5, 1.0, False, False, 4, 8, 2.0).encode(img)
    leaves, so = forest_to_leaves(q, True)
    order = leaf_visiting_order(leaves, W, H)
    nodes, _, n_roots = q.flat_stream()
    want_rects, _ = orc.qt_visit_order(nodes, n_roots, H, W)
    assert np.array_equal(leaves[order][:, :3], want_rects)          # same visiting order as the oracle's walk
    got = []
    dec = QuadtreeDecoder(iterations=3, on_domain_decoded=lambda im, k: got.append((k, im.copy())))
    out = dec.decode(q)
    want = _reference_style_decode_with_callback(leaves[order], so[order], H, W, 3)
    assert [k for k, _ in got] == list(range(1, 3 * len(leaves) + 1))
    for (_, a), b in zip(got, want):
        assert np.array_equal(a, b)
    assert np.array_equal(out, QuadtreeDecoder(iterations=3).decode(q))


def test_auto_stop_follows_the_oracle_across_the_threshold(fic, orc):
    """The AUTO rule stops when ||next - img||_F / ||next||_F < stop_on_relative_error (decoder.py:45-50).  The device uses
    a two-stage deterministic float64 reduction where the reference calls np.linalg.norm: thresholds a hair either side
    of every ratio the decode actually produces must give the same number of sweeps as the oracle."""
    from quadtree.common import forest_to_leaves
    from quadtree.encoder import QuadtreeEncoder
    img = synthetic_image(44, 96, 80)
    q = QuadtreeEncoder(7, 5, 1.0, False, False, 4, 8, 2.0).encode(img)
    leaves, so = forest_to_leaves(q, True)
    H, W = img.shape
    states = [np.zeros((H, W))] + [fic.tree_decode(fic.TREE_QUADTREE, leaves, so, H, W, k, out_f64=True)[0] for k in range(1, 11)]
    ratios = [np.linalg.norm(states[k] - states[k - 1]) / np.linalg.norm(states[k]) for k in range(3, 11)]
    thresholds = [5e-3]
    for r in ratios:
        thresholds += [r * (1 - 1e-6), r * (1 + 1e-6), r * (1 - 1e-10), r * (1 + 1e-10)]
    seen = set()
    for thr in thresholds:
        got, n_got = fic.tree_decode(fic.TREE_QUADTREE, leaves, so, H, W, fic.ITER_AUTO, stop_on_relative_error=thr)
        ref, n_ref = orc.tree_decode(0, leaves, so, H, W, iterations=-1, stop_rel_err=thr)
        assert n_got == n_ref, (thr, n_got, n_ref)
        assert np.array_equal(got, ref)
        seen.add(n_got)
    assert len(seen) >= 3          # the thresholds really straddle several stopping points


def test_quadtree_encoder_keeps_the_reference_attributes(fic, orc):
    """`min_needed_range_exponent` and `domain_lists` (encoder.py:65-75, 261-283) on a ragged image."""
    from quadtree.encoder import QuadtreeEncoder
    img = synthetic_image(7, 44, 36)
    enc = QuadtreeEncoder(7, 5, 1.0, False, False, 4, 8, 2.0)
    enc.encode(img)
    assert enc.max_square_exponent == 5 and enc.min_needed_range_exponent == 2
    lists = enc.domain_lists
    assert len(lists) == enc.max_range_exponent - enc.min_needed_range_exponent + 1
    for k, pool in enumerate(lists):
        side = 2 ** (enc.min_needed_range_exponent + 1 + k)
        assert len(pool) == 3 and all(len(row) == 24 for row in pool)
        total = 0
        for mc in range(3):
            for vc in range(24):
                rows = pool[mc][vc]
                if len(rows) == 1 and rows[0].tolist() == [0, 0, 0, 0]:
                    continue        # placeholder (or the genuine first domain, classified below)
                for (i, j, rot, orient) in rows.tolist():
                    assert tuple(orc.qt_classify(img, i, j, side)) == (mc, vc, rot, orient)
                total += len(rows)
        assert total in (len(range(0, 44 - side + 1, side)) * len(range(0, 36 - side + 1, side)),
                         len(range(0, 44 - side + 1, side)) * len(range(0, 36 - side + 1, side)) - 1)


@pytest.mark.parametrize("params,shape,seed", [
    ((7, 5, 1.0, True, False, 4, 8, 2.0, 1, 1), (128, 128), 51),       # the stress row of DESIGN section 6, small
    ((6, 6, 1.0, True, True, 2, 16, 1.0, 1, 1), (96, 144), 52),        # positive scales only; sizes 2 and 16 stay on CUDA cores
    ((7, 5, 0.75, True, False, 4, 8, 0.5, 1, 1), (104, 88), 53),       # ragged image, smaller max_scale, low tolerance
], ids=["default_bits", "positive_only_2_to_16", "ragged_low_tolerance"])
def test_quadtree_dense_full_class_search_on_the_tensor_cores(fic, orc, params, shape, seed):
    """Dense pools (domain steps 1) searched over all classes go through the tcgen05 screen for 8x8 and 4x4 ranges
    (bf_screen.cu: qt_tc_search); node for node against the oracle, flat and saturated regions included."""
    img = synthetic_image(seed, *shape)
    img[:24, :40] = 255                 # constant blocks: every candidate ties, the scan order decides
    img[40:56, 8:24] = (np.indices((16, 16)).sum(0) % 2 * 255).astype(np.uint8)
    want = orc.qt_encode(img, *params)
    got = fic.qt_encode(img, *params)
    assert got[2] == want[2]
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[1], want[1])


@pytest.mark.parametrize("case", ["synthetic272_seed61_positive", "lena512_stress_row"])
def test_quadtree_dense_full_class_golden(fic, case):
    """The stress row at sizes where the seed pass of the tensor-core search (every 61st tile) is active: SHA-256 of the
    node table, values and root count against the oracle's, computed once on the CPU (tests/golden/make_golden_qt_dense.py)."""
    import hashlib
    import json
    import os
    golden = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "qt_dense_full_golden.json")))
    if case not in golden:
        pytest.skip("no golden for this case (the oracle needs tens of minutes for it)")
    g = golden[case]
    if case.startswith("lena512"):
        img = np.load(os.path.join(os.path.dirname(__file__), "golden", "images.npz"))["lena512"]
    else:
        img = synthetic_image(61, 272, 272)
    assert list(img.shape) == g["shape"]
    nodes, vals, n_roots = fic.qt_encode(img, *g["params"])
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(nodes, dtype=np.int32).tobytes())
    h.update(np.ascontiguousarray(vals, dtype=np.float64).tobytes())
    h.update(int(n_roots).to_bytes(4, "little"))
    assert len(nodes) == g["nodes"]
    assert h.hexdigest() == g["sha256"]


@pytest.mark.parametrize("params,shape,seed", [
    ((7, 5, 1.0, True, False, 4, 8, 2.0), (256, 256), 71),             # default steps (= domain side): 32 x 32 domains for the 4x4 ranges
    ((7, 5, 1.0, True, False, 4, 8, 1.0, 3, 5), (200, 232), 72),       # odd steps, ragged image
    ((6, 6, 0.5, True, True, 4, 16, 3.0, 2, 2), (160, 192), 73),       # positive scales only, smaller max_scale
], ids=["default_steps", "steps_3_5", "steps_2_2_positive"])
def test_quadtree_full_class_search_sparse_pools_on_the_tensor_cores(fic, orc, params, shape, seed):
    """full_class_search with pools that are not dense: the same tensor-core pass over the pool's lattice (qt_tc_search),
    including the placeholder rows of classes that have no domain (quadtree/encoder.py:279-280) -- a flat corner makes
    sure some classes are empty and that constant ranges go to the exact search."""
    img = synthetic_image(seed, *shape)
    img[-40:, -56:] = 17
    want = orc.qt_encode(img, *params)
    got = fic.qt_encode(img, *params)
    assert got[2] == want[2]
    assert np.array_equal(got[0], want[0])
    assert np.array_equal(got[1], want[1])
