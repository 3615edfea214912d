"""GPU: `fic_bf_encode_static`, the entry point for callers that may not allocate device memory (a plugin under
someone else's pool allocator, a serving loop that must stay off cudaMalloc/cudaFree).  It keeps its scratch in
a fixed array inside the library, so the things to check are: same records as the allocating entry point
(`fic_bf_encode`, bit for bit, as everywhere in this suite), still the same on a second call that finds the first
call's leftovers in the scratch, and a clean refusal above the size limit.

The file name sorts last on purpose: a `pytest -x` run has been through every other test before it gets here.
It also passes on its own in a fresh process (nothing here relies on state left by other tests)."""
import numpy as np
import pytest

from conftest import synthetic_image

pytestmark = pytest.mark.gpu


def _run_static(fic, torch, img, mode, strides=(1, 1), begin=0, end=None):
    H, W = img.shape
    total = (H // 8) * (W // 8)
    end = total if end is None else end
    n = end - begin
    d_img = torch.from_numpy(img).cuda()
    d_idx = torch.full((n, 4), -7, dtype=torch.int32, device="cuda")
    d_so = torch.full((n, 2), float("nan"), dtype=torch.float64, device="cuda")
    d_R = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    fic.bf_encode_static_dev(d_img.data_ptr(), H, W, d_idx.data_ptr(), d_so.data_ptr(), d_R.data_ptr(),
                             strides=strides, mode=mode, range_begin=begin, range_end=end)
    torch.cuda.synchronize()
    return d_idx.cpu().numpy(), d_so.cpu().numpy(), d_R.cpu().numpy()


def test_static_entry_matches_allocating_entry_on_two_calls_in_a_row(fic):
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    # call 1: the largest image the built-in scratch takes; call 2: a smaller, different image that finds call 1's
    # data still sitting in the scratch; call 3: a range slice with NumPy semantics and strides
    first = synthetic_image(31, 256, 256)
    second = synthetic_image(32, 64, 96)
    second[:16, :32] = 255
    for img, mode in ((first, fic.BF_MODE_JIT), (second, fic.BF_MODE_JIT)):
        got = _run_static(fic, torch, img, mode)
        want = fic.bf_encode(img, mode=mode, flags=fic.BF_EXACT_ONLY)
        for g, w in zip(got, want):
            assert np.array_equal(g, w)
    got = _run_static(fic, torch, second, fic.BF_MODE_NUMPY, strides=(2, 3), begin=5, end=61)
    want = fic.bf_encode(second, strides=(2, 3), mode=fic.BF_MODE_NUMPY, range_begin=5, range_end=61)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)


def test_static_entry_rejects_images_beyond_its_scratch(fic):
    torch = pytest.importorskip("torch")
    torch.cuda.set_device(0)
    big = torch.zeros((264, 256), dtype=torch.uint8, device="cuda")
    out = torch.zeros(1 << 16, dtype=torch.float64, device="cuda")
    with pytest.raises(ValueError, match="exceeds the built-in scratch"):
        fic.bf_encode_static_dev(big.data_ptr(), 264, 256, out.data_ptr(), out.data_ptr(), out.data_ptr())
