"""GPU: the brute-force encode with device pointers does not wait for the GPU (include/fic.h): calls queue up behind
each other, the ranges the tensor-core screen cannot decide are re-scanned by a launch that sizes itself from a
device-side count, the counters are kept on the device until somebody asks, and calls on different streams or
different GPUs of one process do not disturb each other."""
import hashlib
import os

import numpy as np
import pytest

from conftest import GOLDEN, synthetic_image

pytestmark = pytest.mark.gpu


def _records_sha(idx, so, R) -> str:
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(idx, np.int32).tobytes())
    h.update(np.ascontiguousarray(so, np.float64).tobytes())
    h.update(np.ascontiguousarray(R, np.float64).tobytes())
    return h.hexdigest()


def _dev_encode(fic, torch, h, d_img, H, W, stream=0, begin=0, end=None, mode=1):
    n = (H // 8) * (W // 8) if end is None else end - begin
    d_idx = torch.full((n, 4), -3, dtype=torch.int32, device=d_img.device)
    d_so = torch.full((n, 2), float("nan"), dtype=torch.float64, device=d_img.device)
    d_R = torch.full((n,), float("nan"), dtype=torch.float64, device=d_img.device)
    fic.bf_encode_dev(h, d_img.data_ptr(), H, W, d_idx.data_ptr(), d_so.data_ptr(), d_R.data_ptr(), mode=mode,
                      range_begin=begin, range_end=end, stream=stream)
    return d_idx, d_so, d_R


def test_lena512_all_4096_records_equal_the_oracle_sha(fic, images):
    """The benchmark's image, every record, against the CPU oracle's full encode (tests/golden/make_golden_full.py)."""
    g = np.load(os.path.join(GOLDEN, "bf_full_golden.npz"))
    idx, so, R = fic.bf_encode(images["lena512"], mode=fic.BF_MODE_JIT)
    assert np.array_equal(idx, g["jit512_idx"])
    assert _records_sha(idx, so, R) == str(g["jit512_sha256"])


def test_synthetic_4096_ranges_equal_the_oracle(fic):
    """configs[3] at full size: 16 ranges of the synthetic 4096x4096 image against the CPU oracle's records."""
    g = np.load(os.path.join(GOLDEN, "bf_full_golden.npz"))
    img = synthetic_image(0, 4096, 4096)
    assert hashlib.sha256(img.tobytes()).hexdigest() == str(g["syn4096_image_sha256"])
    for k, rec in zip(g["syn4096_sel"], g["syn4096_rec"]):
        # a slice of 32 ranges around the golden one: one column group of the screen
        b = int(k) // 32 * 32
        idx, so, R = fic.bf_encode(img, mode=fic.BF_MODE_JIT, range_begin=b, range_end=b + 32)
        j = int(k) - b
        assert idx[j].tolist() == [int(v) for v in rec[:4]]
        assert so[j, 0] == rec[4] and so[j, 1] == rec[5] and R[j] == rec[6]


def test_device_pointer_calls_queue_without_waiting_and_agree(fic, images):
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    h = fic.default_handle(0)
    img = np.ascontiguousarray(images["lena256"])
    want = fic.bf_encode(img, mode=1, flags=fic.BF_EXACT_ONLY, handle=h)
    d_img = torch.from_numpy(img).cuda()
    st = torch.cuda.Stream()
    h.reset_counters()
    outs = [_dev_encode(fic, torch, h, d_img, 256, 256, stream=st.cuda_stream) for _ in range(6)]
    st.synchronize()
    for d_idx, d_so, d_R in outs:
        assert np.array_equal(d_idx.cpu().numpy(), want[0])
        assert np.array_equal(d_so.cpu().numpy(), want[1])
        assert np.array_equal(d_R.cpu().numpy(), want[2])
    c = h.counters()
    assert c["screen_ranges"] + c["rescan_ranges"] == 6 * 1024 and c["screen_ranges"] > 0


def _impulse_lattice(H, W, at=(3, 5)):
    """255 everywhere except one 254 per 8x8 range block, at the same place in every block.  Every 16x16 window then
    holds exactly four impulses, so the best candidate explains a quarter of a range's (tiny) variance -- less than the
    float32 noise allowance of the screen's bound (64e-9 * sum r^2 = 0.27 for a bright block): no range can be screened."""
    img = np.full((H, W), 255, np.uint8)
    img[at[0]::8, at[1]::8] = 254
    return img


def test_device_side_rescan_of_undecided_ranges(fic):
    """Ranges the screen cannot decide are re-scanned by a launch that takes their number from the device: all of
    them (more range groups than the launch has CTAs), some of them, and a second call that must find the arrival
    counters of the first one back at zero."""
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    h = fic.default_handle(0)
    img = _impulse_lattice(160, 192)
    n = 20 * 24
    want = fic.bf_encode(img, mode=1, flags=fic.BF_EXACT_ONLY, handle=h)
    d_img = torch.from_numpy(img).cuda()
    h.reset_counters()
    got = _dev_encode(fic, torch, h, d_img, 160, 192)
    again = _dev_encode(fic, torch, h, d_img, 160, 192)
    torch.cuda.synchronize()
    for g, a, w in zip(got, again, want):
        assert np.array_equal(g.cpu().numpy(), w)
        assert np.array_equal(a.cpu().numpy(), w)
    c = h.counters()
    assert c["rescan_ranges"] == 2 * n and c["screen_ranges"] == 0, c
    # half lattice, half texture; NumPy semantics on a slice; host buffers
    img[80:] = synthetic_image(11, 80, 192)
    for mode, (b, e) in ((1, (0, n)), (0, (100, 333))):
        want = fic.bf_encode(img, mode=mode, flags=fic.BF_EXACT_ONLY, range_begin=b, range_end=e, handle=h)
        h.reset_counters()
        host = fic.bf_encode(img, mode=mode, range_begin=b, range_end=e, handle=h)
        c = h.counters()
        assert 0 < c["rescan_ranges"] < e - b and c["screen_ranges"] + c["rescan_ranges"] == e - b, c
        got = _dev_encode(fic, torch, h, torch.from_numpy(img).cuda(), 160, 192, begin=b, end=e, mode=mode)
        torch.cuda.synchronize()
        for g, hst, w in zip(got, host, want):
            assert np.array_equal(g.cpu().numpy(), w)
            assert np.array_equal(hst, w)
    # the testing flag that refuses to re-scan reports the undecided ranges instead
    with pytest.raises(fic.FicError, match="undecided"):
        fic.bf_encode(img, mode=1, flags=fic.BF_NO_FALLBACK, handle=h)


def test_calls_on_two_streams_are_ordered_by_the_handle(fic, images):
    """Two device-pointer calls on different streams share the handle's workspaces: the second is made to wait."""
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    h = fic.default_handle(0)
    a_img = np.ascontiguousarray(images["lena256"][:128, :128])
    b_img = synthetic_image(5, 128, 128)
    want_a = fic.bf_encode(a_img, mode=1, handle=h)
    want_b = fic.bf_encode(b_img, mode=1, handle=h)
    d_a, d_b = torch.from_numpy(a_img).cuda(), torch.from_numpy(b_img).cuda()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    outs = []
    for rep in range(4):
        outs.append((_dev_encode(fic, torch, h, d_a, 128, 128, stream=s1.cuda_stream), want_a))
        outs.append((_dev_encode(fic, torch, h, d_b, 128, 128, stream=s2.cuda_stream), want_b))
    torch.cuda.synchronize()
    for got, want in outs:
        for g, w in zip(got, want):
            assert np.array_equal(g.cpu().numpy(), w)


def test_caller_device_is_left_alone(fic):
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    torch.cuda.set_device(1)
    h0 = fic.default_handle(0)
    img = synthetic_image(3, 96, 96)
    fic.bf_encode(img, handle=h0)
    assert torch.cuda.current_device() == 1
    torch.cuda.set_device(0)


def test_handles_on_two_gpus_in_one_process(fic):
    """The screen kernels opt in to > 48 KB of dynamic shared memory per device: a second GPU needs its own opt-in."""
    torch = pytest.importorskip("torch")
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    img = synthetic_image(9, 96, 128)
    h0, h1 = fic.default_handle(0), fic.Handle(1)
    try:
        a = fic.bf_encode(img, handle=h0)
        h1.reset_counters()
        b = fic.bf_encode(img, handle=h1)
        assert h1.counters()["screen_ranges"] > 0
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
    finally:
        h1.close()
