"""GPU: brute-force encoders and decoder through the C ABI (via the reference-shaped classes) against the
reference's golden outputs and against the CPU oracle on seeded inputs.  Integer/index work and the float64
(s, o, R) values are compared BIT-EXACT."""
import numpy as np
import pytest

from conftest import synthetic_image

pytestmark = pytest.mark.gpu


def records(enc):
    idx = np.array([[d.start_i, d.start_j, d.orientation, d.rotation] for d in enc.domains], np.int32)
    so = np.array([[d.s, d.o] for d in enc.domains], np.float64)
    return idx, so


@pytest.mark.parametrize("tag", ["jit48", "jitflat", "jitchk"])
def test_jit_encoder_matches_reference_golden(fic, golden, tag):
    from bruteforce.jit_bfencoder import JitedBruteForceEncoder
    g = golden["bf_golden"]
    enc, errs = JitedBruteForceEncoder(g[f"{tag}_img"]).encode()
    idx, so = records(enc)
    assert np.array_equal(idx, g[f"{tag}_idx"])
    assert np.array_equal(so, g[f"{tag}_so"])
    assert np.array_equal(np.array(errs), g[f"{tag}_errs"])
    assert (enc.width, enc.height, enc.range_size) == (48, 48, 8)


@pytest.mark.parametrize("flags", [0, 2])  # default pipeline and FIC_BF_EXACT_ONLY
def test_jit_full_256_matches_reference(fic, golden, images, flags):
    g = golden["bf_golden"]
    idx, so, R = fic.bf_encode(images["lena256"], mode=fic.BF_MODE_JIT, flags=flags)
    assert np.array_equal(idx, g["jit256_idx"])
    assert np.array_equal(so, g["jit256_so"])
    assert np.array_equal(np.sqrt(R), g["jit256_errs"])


@pytest.mark.parametrize("flags", [0, 2])
def test_jit_512_selected_ranges_match_reference(fic, golden, images, flags):
    g = golden["bf_golden"]
    idx, so, R = fic.bf_encode(images["lena512"], mode=fic.BF_MODE_JIT, flags=flags)
    for k, rec in zip(g["jit512_sel"], g["jit512_rec"]):
        assert idx[k].tolist() == [int(v) for v in rec[:4]]
        assert so[k, 0] == rec[4] and so[k, 1] == rec[5] and R[k] == rec[6]


def test_pipeline_equals_exact_scan_on_512(fic, images):
    """Size-independent property at the headline size: the screened pipeline and the exhaustive exact scan
    must agree record for record."""
    a = fic.bf_encode(images["lena512"], mode=fic.BF_MODE_JIT, flags=0)
    b = fic.bf_encode(images["lena512"], mode=fic.BF_MODE_JIT, flags=fic.BF_EXACT_ONLY)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_numpy_encoder_matches_reference_golden(fic, golden):
    from bruteforce.bf_encoder import BruteForceEncoder
    g = golden["bf_golden"]
    idx, so = records(BruteForceEncoder(g["np32_img"]).encode())
    assert np.array_equal(idx, g["np32_idx"]) and np.array_equal(so, g["np32_so"])
    idx, so = records(BruteForceEncoder(g["np32_img"], strides=(2, 3)).encode())
    assert np.array_equal(idx, g["np32s23_idx"]) and np.array_equal(so, g["np32s23_so"])


@pytest.mark.parametrize("seed,shape", [(1, (64, 64)), (2, (48, 80)), (3, (96, 32))])
@pytest.mark.parametrize("mode", [0, 1])
def test_against_oracle_on_seeded_images(fic, orc, seed, shape, mode):
    img = synthetic_image(seed, *shape)
    if seed == 3:
        img[:24] = 255  # saturated region: flat ranges and flat domains, ties decided by scan order
        img[40:56, 8:24] = 0
    want = orc.bf_encode(img, mode=mode)
    got = fic.bf_encode(img, mode=mode)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)
    if mode == 1:
        got = fic.bf_encode(img, mode=mode, flags=fic.BF_EXACT_ONLY)
        for x, y in zip(got, want):
            assert np.array_equal(x, y)


@pytest.mark.parametrize("rs", [4, 16])
@pytest.mark.parametrize("mode", [0, 1])
def test_other_range_sizes_against_oracle(fic, orc, rs, mode):
    img = synthetic_image(5, 64, 64)
    want = orc.bf_encode(img, range_size=rs, domain_size=2 * rs, mode=mode)
    got = fic.bf_encode(img, range_size=rs, domain_size=2 * rs, mode=mode)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)


@pytest.mark.parametrize("shape,strides,seed", [((96, 96), (1, 1), 21), ((128, 100), (2, 1), 22), ((104, 120), (1, 3), 23)],
                         ids=["96x96", "128x100_strides_2_1", "104x120_strides_1_3"])
@pytest.mark.parametrize("mode", [1, 0], ids=["jit", "numpy"])
def test_range_size_4_on_the_tensor_cores(fic, orc, shape, strides, seed, mode):
    """4x4 ranges (both encoders' arithmetic: every sum is exact with sixteen pixels) go through the K = 16 tensor-core pass (bf_screen.cu: bf4_tc_encode) once
    the lattice has 4096 positions; a flat patch (constant ranges: handed to the exact scan), a saturated one and a
    checkerboard are in the image.  All records against the oracle, and against the library's own exact scan."""
    img = synthetic_image(seed, *shape)
    img[:12, :20] = 200
    img[40:52, 60:72] = 255
    img[60:68, 8:16] = (np.indices((8, 8)).sum(0) % 2 * 255).astype(np.uint8)
    want = orc.bf_encode(img, range_size=4, domain_size=8, strides=strides, mode=mode)
    got = fic.bf_encode(img, range_size=4, domain_size=8, strides=strides, mode=mode)
    exact = fic.bf_encode(img, range_size=4, domain_size=8, strides=strides, mode=mode, flags=fic.BF_EXACT_ONLY)
    for x, y, z in zip(got, want, exact):
        assert np.array_equal(x, y)
        assert np.array_equal(z, y)


@pytest.mark.parametrize("rs,shape,strides,seed", [(16, (80, 112), (1, 1), 31), (16, (96, 96), (3, 2), 32), (32, (128, 160), (1, 1), 33),
                                                   (2, (40, 36), (1, 1), 34)],
                         ids=["r16_35_ranges", "r16_strides_3_2", "r32", "r2"])
@pytest.mark.parametrize("mode", [1, 0], ids=["jit", "numpy"])
def test_generic_range_sizes_four_ranges_per_cta(fic, orc, rs, shape, strides, seed, mode):
    """Range sizes other than 4 and 8 (both encoders' arithmetic): four ranges per CTA over per-position tables of
    sum d and sum d^2 (bf_exact.cu: bf_scan_generic4_kernel), range counts that are not a multiple of four included."""
    img = synthetic_image(seed, *shape)
    img[:rs, :rs] = 128
    want = orc.bf_encode(img, range_size=rs, domain_size=2 * rs, strides=strides, mode=mode)
    got = fic.bf_encode(img, range_size=rs, domain_size=2 * rs, strides=strides, mode=mode)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)


def test_random_noise_and_extremes_against_oracle(fic, orc):
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (48, 48), dtype=np.uint8)
    img[:8, :8] = 255
    img[8:16, :8] = 0
    for mode in (0, 1):
        want = orc.bf_encode(img, mode=mode)
        got = fic.bf_encode(img, mode=mode)
        for x, y in zip(got, want):
            assert np.array_equal(x, y)


def test_constant_image_ties_follow_scan_order(fic, orc):
    img = np.full((48, 48), 137, np.uint8)
    idx, so, R = fic.bf_encode(img, mode=1)
    want = orc.bf_encode(img, mode=1)
    assert np.array_equal(idx, want[0]) and np.array_equal(so, want[1]) and np.array_equal(R, want[2])
    assert (idx == 0).all()  # first candidate in scan order wins every tie


def test_range_slices_concatenate_to_the_full_encode(fic, images):
    """Sharding by range-block slices needs no exchange: slices are independent."""
    img = images["lena256"][:128, :128]
    img = np.ascontiguousarray(img)
    full = fic.bf_encode(img, mode=1)
    parts = [fic.bf_encode(img, mode=1, range_begin=b, range_end=e) for b, e in ((0, 100), (100, 101), (101, 256))]
    for k in range(3):
        assert np.array_equal(np.concatenate([p[k] for p in parts]), full[k])
    empty = fic.bf_encode(img, mode=1, range_begin=7, range_end=7)
    assert empty[0].shape == (0, 4)


def test_image_without_any_domain_position(fic, orc):
    img = synthetic_image(9, 16, 16)
    got = fic.bf_encode(img, mode=1)
    want = orc.bf_encode(img, mode=1)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)
    from bruteforce.bf_encoder import BruteForceEncoder
    with pytest.raises(TypeError):
        BruteForceEncoder(img).encode()


def test_bad_arguments_raise(fic):
    img = synthetic_image(1, 64, 64)
    with pytest.raises(ValueError):
        fic.bf_encode(img, range_size=8, domain_size=12)
    with pytest.raises(ValueError):
        fic.bf_encode(img[:60], mode=1)
    with pytest.raises(ValueError):
        fic.bf_encode(img, mode=1, range_begin=5, range_end=1000)
    with pytest.raises(TypeError):
        fic.bf_encode(img.astype(np.float32))


def test_decoder_matches_reference_golden(fic, golden):
    from bruteforce.decoder import Decoder
    from bruteforce.model import EncodedImage
    g = golden["bf_golden"]
    for tag, n in (("jit48", 48), ("jitflat", 48), ("np32", 32)):
        enc = EncodedImage.from_arrays(n, n, 8, g[f"{tag}_idx"], g[f"{tag}_so"])
        assert np.array_equal(Decoder().decode(enc, 32), g[f"{tag}_dec32"])
    enc = EncodedImage.from_arrays(48, 48, 8, g["jit48_idx"], g["jit48_so"])
    assert np.array_equal(Decoder().decode(enc, 3), g["jit48_dec3"])
    assert np.array_equal(Decoder().decode(enc, 0), np.zeros((48, 48), np.uint8))
    enc = EncodedImage.from_arrays(256, 256, 8, g["jit256_idx"], g["jit256_so"])
    dec = Decoder().decode(enc)
    assert np.array_equal(dec, g["jit256_dec32"])


def test_decoder_against_oracle_unbounded_scale(fic, orc):
    """NumPy-mode transforms are not contractive (|s| unbounded): values leave [0,255] and the uint8 cast wraps."""
    img = synthetic_image(4, 64, 64)
    idx, so, _ = orc.bf_encode(img, mode=0)
    for it in (1, 2, 7, 32):
        assert np.array_equal(fic.bf_decode(idx, so, 64, 64, it), orc.bf_decode(idx, so, 64, 64, it))


def test_decoder_values_beyond_int16_take_the_global_memory_path(fic, orc):
    """Small images are decoded against an int16 copy of the truncated pixels in shared memory; transforms that push
    pixels past +-32767 must fall back (inside the same launch) to the float image in global memory and still agree with
    the oracle.  One plane of the batch overflows, its neighbours do not."""
    rng = np.random.default_rng(5)
    n = 64
    idx = np.stack([np.c_[rng.integers(0, 49, (n, 2)), rng.integers(0, 2, n), rng.integers(0, 4, n)] for _ in range(3)]).astype(np.int32)
    so = np.zeros((3, n, 2))
    so[0] = (0.6, 40.0)
    so[1] = (2.0, 10.0)            # doubles every sweep: ~2.6e5 after ten sweeps
    so[2] = (-0.9, 200.0)
    for it in (1, 5, 9, 12):
        got = fic.bf_decode_batch(idx, so, 64, 64, it)
        for k in range(3):
            assert np.array_equal(got[k], orc.bf_decode(idx[k], so[k], 64, 64, it)), (it, k)
    # the image just above the shared-memory limit (328 x 328 > 102 400 pixels) runs the global-memory kernel from the start
    img = synthetic_image(6, 328, 328)
    e = fic.bf_encode(img, mode=1)
    assert np.array_equal(fic.bf_decode(e[0], e[1], 328, 328, 6), orc.bf_decode(e[0], e[1], 328, 328, 6))
    img = synthetic_image(7, 320, 320)
    e = fic.bf_encode(img, mode=1)
    assert np.array_equal(fic.bf_decode(e[0], e[1], 320, 320, 6), orc.bf_decode(e[0], e[1], 320, 320, 6))


@pytest.mark.parametrize("shape", [(512, 512), (328, 328), (1024, 1024), (776, 776)])
def test_decoder_large_images_match_oracle(fic, orc, shape):
    """Images above the shared-memory limit of 102 400 pixels (global-memory path): random contractive records, records
    of a real encode, and non-contractive records next to contractive ones in one batch, against the oracle."""
    H, W = shape
    img = synthetic_image(9, H, W)
    n = (H // 8) * (W // 8)
    rng = np.random.default_rng(H)
    idx = np.c_[rng.integers(0, H - 15, n), rng.integers(0, W - 15, n), rng.integers(0, 2, n), rng.integers(0, 4, n)].astype(np.int32)
    so = np.c_[rng.uniform(-0.9, 0.9, n), rng.uniform(0, 255, n)]
    for it in (1, 2, 5):
        assert np.array_equal(fic.bf_decode(idx, so, W, H, it), orc.bf_decode(idx, so, W, H, it)), it
    if H <= 512:
        e = fic.bf_encode(img, mode=1)
        assert np.array_equal(fic.bf_decode(e[0], e[1], W, H, 8), orc.bf_decode(e[0], e[1], W, H, 8))
    so2 = so.copy()
    so2[:, 0] = 2.0
    both = fic.bf_decode_batch(np.stack([idx, idx]), np.stack([so, so2]), W, H, 11)
    assert np.array_equal(both[0], orc.bf_decode(idx, so, W, H, 11))
    assert np.array_equal(both[1], orc.bf_decode(idx, so2, W, H, 11))


def test_encode_decode_round_trip_psnr(fic, images, golden):
    """PSNR of GPU encode -> GPU decode equals the reference's 29.967 dB on lena 256."""
    img = images["lena256"]
    idx, so, _ = fic.bf_encode(img, mode=1)
    dec = fic.bf_decode(idx, so, 256, 256, 32)
    mse = np.mean((dec.astype(np.float64) - img.astype(np.float64)) ** 2)
    psnr = 10 * np.log10(255.0 ** 2 / mse)
    assert abs(psnr - 29.96704623) < 1e-6


def test_batch_decode_planes_run_concurrently_and_match(fic, orc):
    """fic_bf_decode_batch: independent planes in one launch, each equal to its own single decode and to the oracle;
    mixes contractive (JIT) and non-contractive (NumPy mode) transforms and an odd number of sweeps."""
    imgs = [synthetic_image(30 + k, 64, 64) for k in range(5)]
    recs = [orc.bf_encode(im, mode=k % 2) for k, im in enumerate(imgs)]
    idx = np.stack([r[0] for r in recs]); so = np.stack([r[1] for r in recs])
    for it in (1, 2, 9):
        got = fic.bf_decode_batch(idx, so, 64, 64, it)
        for k in range(5):
            assert np.array_equal(got[k], orc.bf_decode(idx[k], so[k], 64, 64, it))
            assert np.array_equal(got[k], fic.bf_decode(idx[k], so[k], 64, 64, it))
    assert fic.bf_decode_batch(idx[:0], so[:0], 64, 64, 3).shape == (0, 64, 64)
    with pytest.raises(ValueError):
        fic.bf_decode_batch(idx, so[:, :10], 64, 64, 3)


def test_batch_entry_point(fic, orc):
    imgs = np.stack([synthetic_image(20 + k, 48, 48) for k in range(3)])
    idx, so, R = fic.bf_encode_batch(imgs)
    for k in range(3):
        want = orc.bf_encode(imgs[k], mode=1)
        assert np.array_equal(idx[k], want[0]) and np.array_equal(so[k], want[1]) and np.array_equal(R[k], want[2])


def test_average_subsample(fic, orc):
    import utils
    rng = np.random.default_rng(3)
    a = rng.random((16, 12)) * 300 - 20
    assert np.array_equal(utils.average_subsample(a), orc.average_subsample(a, jit=False))
    got = utils.average_subsample_jit(a)
    assert got.dtype == np.float32
    assert np.array_equal(got.astype(np.float64), orc.average_subsample(a, jit=True))
    with pytest.raises(ValueError):
        fic.average_subsample(np.zeros((3, 4)))
