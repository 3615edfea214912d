"""GPU: the tensor-core screen of the brute-force encoder.

(1) `fic_bf_screen_probe` returns the raw tcgen05 accumulators; they must equal the integer X' = sum r'_k D4'_k
    computed here (this pins the TMA tensor maps, the UMMA shared-memory and instruction descriptors, the TMEM
    lane/column mapping and the fp32 accumulation all at once), and the reference's numerator u = X' - rho*eps/64 must
    lie within the DELTA_U = 16 the screen's thresholds allow for.
(2) The full pipeline (pool -> screen -> exact rescoring [-> exact re-scan]) must return the same records as the
    exhaustive exact scan, bit for bit, on natural, synthetic, saturated and flat-heavy images.
"""
import numpy as np
import pytest

from conftest import synthetic_image

pytestmark = pytest.mark.gpu


def exact_u(img, range_begin, n_ranges, with_rank1=False):
    img = img.astype(np.int64)
    H, W = img.shape
    box = img[:-1, :-1] + img[:-1, 1:] + img[1:, :-1] + img[1:, 1:]          # P[y][x]
    PH, PW = H - 16, W - 16
    # D4[p, a, b] = P[di + 2a, dj + 2b]
    D4 = np.empty((PH, PW, 8, 8), np.int64)
    for a in range(8):
        for b in range(8):
            D4[:, :, a, b] = box[2 * a:2 * a + PH, 2 * b:2 * b + PW]
    D4 = D4.reshape(PH * PW, 8, 8)
    sd4 = D4.sum((1, 2))
    m_c = (sd4 + 32) >> 6
    eps = sd4 - 64 * m_c
    Dc = D4 - m_c[:, None, None]
    rw = W // 8
    out = np.empty((PH * PW, n_ranges * 8), np.float64)
    for k in range(n_ranges):
        ridx = range_begin + k
        ri, rj = (ridx // rw) * 8, (ridx % rw) * 8
        r = img[ri:ri + 8, rj:rj + 8]
        s = int(r.sum())
        m_r = (s + 32) >> 6
        rho = s - 64 * m_r
        rc = r - m_r
        for orient in range(2):
            base = Dc.transpose(0, 2, 1) if orient else Dc
            for rot in range(4):
                Y = np.rot90(base, rot, axes=(1, 2))                             # jit_bfencoder.py:44-49
                out[:, k * 8 + orient * 4 + rot] = (Y * rc[None]).sum((1, 2)) - (rho * eps / 64.0 if with_rank1 else 0.0)
    return out


@pytest.mark.parametrize("case", ["synthetic", "extremes", "lena_offset"])
def test_probe_accumulators_are_exact(fic, images, case):
    if case == "synthetic":
        img, begin, n = synthetic_image(3, 96, 96), 0, 32
    elif case == "extremes":
        rng = np.random.default_rng(5)
        img = (np.indices((96, 112)).sum(0) % 2 * 255).astype(np.uint8)   # 0/255 checkerboard: largest products
        img[:32] = 255
        img[32:48, :48] = 0
        img[60:, 50:] = rng.integers(0, 256, (36, 62))
        begin, n = 7, 32
    else:
        img, begin, n = np.ascontiguousarray(images["lena256"][:128, :160]), 100, 17
    got = fic.bf_screen_probe(img, begin, n).astype(np.float64)
    want = exact_u(img, begin, n)
    # integers below 2^24: the fp32 accumulator is exact, always
    assert np.array_equal(got, want), f"max |X_tc - X| = {np.abs(got - want).max()}"
    # ... and the rank-1 centring term it leaves out never exceeds the slack of the thresholds
    assert np.abs(got - exact_u(img, begin, n, with_rank1=True)).max() <= 16.0


@pytest.mark.parametrize("mode", [1, 0], ids=["jit", "numpy"])
@pytest.mark.parametrize("seed,shape", [(1, (96, 96)), (2, (128, 200)), (3, (160, 96))])
def test_pipeline_equals_exact_scan(fic, seed, shape, mode):
    img = synthetic_image(seed, *shape)
    if seed == 3:
        img[:40] = 255          # saturated: flat ranges the screen must hand to the exact scan
        img[100:120, 10:60] = 0
    h = fic.default_handle()
    h.reset_counters()
    a = fic.bf_encode(img, mode=mode, flags=0)
    c = h.counters()
    b = fic.bf_encode(img, mode=mode, flags=fic.BF_EXACT_ONLY)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    n = (shape[0] // 8) * (shape[1] // 8)
    assert c["screen_ranges"] + c["rescan_ranges"] == n
    if seed != 3:
        assert c["screen_ranges"] > 0


@pytest.mark.parametrize("mode", [1, 0], ids=["jit", "numpy"])
def test_pipeline_noise_image(fic, mode):
    """White noise: every candidate is about equally bad, the survivor lists are long."""
    rng = np.random.default_rng(17)
    img = rng.integers(0, 256, (96, 96), dtype=np.uint8)
    a = fic.bf_encode(img, mode=mode, flags=0)
    b = fic.bf_encode(img, mode=mode, flags=fic.BF_EXACT_ONLY)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_numpy_mode_pipeline_equals_oracle(fic, orc, images):
    """BruteForceEncoder semantics (float64, no clamp, uint8-wrapped sum of squares) through the tensor-core screen:
    the records equal the CPU restatement of bf_encoder.py, which is pinned to the reference in tests/golden."""
    img = np.ascontiguousarray(images["lena256"][96:192, 64:176])       # 96 x 112: 6 400 candidate positions
    h = fic.default_handle()
    h.reset_counters()
    got = fic.bf_encode(img, mode=fic.BF_MODE_NUMPY)
    assert h.counters()["screen_ranges"] > 0
    want = orc.bf_encode(img, mode=0)
    for x, y in zip(got, want):
        assert np.array_equal(x, y)


def test_pipeline_range_slice(fic, images):
    img = np.ascontiguousarray(images["lena256"][64:192, 64:192])
    full = fic.bf_encode(img, mode=fic.BF_MODE_JIT, flags=fic.BF_EXACT_ONLY)
    part = fic.bf_encode(img, mode=fic.BF_MODE_JIT, range_begin=37, range_end=201)
    for x, y in zip(part, full):
        assert np.array_equal(x, y[37:201])


def test_pipeline_counts_on_lena512(fic, images):
    h = fic.default_handle()
    h.reset_counters()
    fic.bf_encode(images["lena512"], mode=fic.BF_MODE_JIT)
    c = h.counters()
    assert c["screen_ranges"] + c["rescan_ranges"] == 4096
    assert c["screen_ranges"] >= 3500, c          # natural image: nearly every range is decided by the screen
    print("lena512 screen counters:", c)


def test_three_round_plan_on_a_large_image(fic):
    """1024 x 1536: 1.5 M candidate positions -> the round plan has an extra sparse round and several L2 windows.
    The full pipeline over all ranges is compared with the exhaustive exact scan on two range slices."""
    img = synthetic_image(41, 1024, 1536)
    img[200:264, 300:420] = 17                 # a constant patch: constant ranges are decided directly
    h = fic.default_handle()
    h.reset_counters()
    idx, so, R = fic.bf_encode(img, mode=fic.BF_MODE_JIT)
    c = h.counters()
    n = (1024 // 8) * (1536 // 8)
    assert c["screen_ranges"] + c["rescan_ranges"] == n
    for b, e in ((0, 192), (5000, 5256), (n - 64, n)):
        want = fic.bf_encode(img, mode=fic.BF_MODE_JIT, flags=fic.BF_EXACT_ONLY, range_begin=b, range_end=e)
        assert np.array_equal(idx[b:e], want[0]) and np.array_equal(so[b:e], want[1]) and np.array_equal(R[b:e], want[2])
    # the same slice through the pipeline on its own (range-slice sharding of a big image)
    part = fic.bf_encode(img, mode=fic.BF_MODE_JIT, range_begin=5000, range_end=5256)
    assert np.array_equal(part[0], idx[5000:5256]) and np.array_equal(part[1], so[5000:5256])


def test_full_size_4096_properties(fic):
    """BASELINE configs[3] at its full size (4096 x 4096, 262 144 ranges x 16.6 M candidates x 8 isometries -- out of any
    CPU oracle's reach): size-independent properties.  Scattered ranges equal the exhaustive exact scan, a 1/8 range slice
    run on its own (what one rank of eight does) equals the same rows of the full encode, every record points inside the
    image, and no range was left undecided."""
    img = synthetic_image(0, 4096, 4096)
    n = 512 * 512
    h = fic.default_handle()
    h.reset_counters()
    idx, so, R = fic.bf_encode(img, mode=fic.BF_MODE_JIT)
    c = h.counters()
    assert c["screen_ranges"] + c["rescan_ranges"] == n
    assert idx[:, 0].min() >= 0 and idx[:, 1].min() >= 0 and idx[:, 0].max() <= 4096 - 16 and idx[:, 1].max() <= 4096 - 16
    assert np.isfinite(so).all() and np.isfinite(R).all() and (np.abs(so[:, 0]) <= 1.0).all() and (R >= 0).all()
    for k in (0, 777, 131071, 200000, n - 1):
        w = fic.bf_encode(img, mode=fic.BF_MODE_JIT, flags=fic.BF_EXACT_ONLY, range_begin=k, range_end=k + 1)
        assert np.array_equal(w[0][0], idx[k]) and np.array_equal(w[1][0], so[k]) and w[2][0] == R[k]
    b, e = 3 * n // 8, 4 * n // 8
    part = fic.bf_encode(img, mode=fic.BF_MODE_JIT, range_begin=b, range_end=e)
    assert np.array_equal(part[0], idx[b:e]) and np.array_equal(part[1], so[b:e]) and np.array_equal(part[2], R[b:e])


def _stress_images():
    rng = np.random.default_rng(2024)
    yy, xx = np.mgrid[0:160, 0:192]
    out = {}
    out["gradient"] = ((yy * 1.3 + xx * 0.4) % 256).astype(np.uint8)                       # smooth ramps: near-flat domains
    out["low_contrast"] = (120 + rng.integers(0, 4, (160, 192))).astype(np.uint8)            # variance below the guard
    out["steps"] = ((xx // 24) * 36 % 256).astype(np.uint8)                                  # piecewise constant: den16 == 0 often
    out["steps_noise"] = np.clip(out["steps"].astype(int) + rng.integers(-1, 2, (160, 192)), 0, 255).astype(np.uint8)
    out["saturated"] = np.where(rng.random((160, 192)) < 0.5, 0, 255).astype(np.uint8)       # largest magnitudes
    out["sparse_dots"] = np.zeros((160, 192), np.uint8)
    out["sparse_dots"][rng.integers(0, 160, 300), rng.integers(0, 192, 300)] = 255
    s = synthetic_image(77, 160, 192).astype(int)
    out["dark"] = (s // 16).astype(np.uint8)                                                 # tiny dynamic range
    out["bright_clip"] = np.clip(s * 2, 0, 255).astype(np.uint8)                             # half the image clipped to 255
    out["texture_tiles"] = np.tile(rng.integers(0, 256, (16, 16), dtype=np.uint8), (10, 12)) # exact repeats: many exact ties
    out["self_similar"] = np.kron(rng.integers(0, 256, (20, 24)), np.ones((8, 8))).astype(np.uint8)  # every range constant
    return out


@pytest.mark.parametrize("mode", [1, 0], ids=["jit", "numpy"])
@pytest.mark.parametrize("name", sorted(_stress_images()))
def test_screen_is_conservative_on_adversarial_images(fic, name, mode):
    """The screen may only drop candidates that cannot win or tie.  Images chosen to attack the bound: domains with
    (almost) zero variance, ranges flatter than the float32 noise guard, saturated pixels, exact repeats (ties)."""
    img = _stress_images()[name]
    a = fic.bf_encode(img, mode=mode, flags=0)
    b = fic.bf_encode(img, mode=mode, flags=fic.BF_EXACT_ONLY)
    for x, y in zip(a, b):
        assert np.array_equal(x, y), name
