import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "fractal-image-compression_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _has_gpu() -> bool:
    # a library that does not load is a build error, not "no GPU": let it propagate so the run fails loudly
    import fic_b200
    return fic_b200.lib().fic_device_count() > 0


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Golden:
    def __init__(self):
        self._cache = {}

    def __getitem__(self, name):
        if name not in self._cache:
            self._cache[name] = np.load(os.path.join(GOLDEN, name + ".npz"))
        return self._cache[name]

    def path(self, fname):
        return os.path.join(GOLDEN, fname)


@pytest.fixture(scope="session")
def golden():
    return Golden()


@pytest.fixture(scope="session")
def images(golden):
    return golden["images"]


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def fic():
    import fic_b200
    fic_b200.default_handle()
    return fic_b200


def synthetic_image(seed: int, H: int, W: int) -> np.ndarray:
    """Four octaves of bilinearly up-sampled uniform noise + 8 % white noise, rescaled to 0..255."""
    rng = np.random.default_rng(seed)
    acc = np.zeros((H, W), np.float64)
    for o, cell in enumerate((64, 32, 16, 8)):
        gh, gw = H // cell + 2, W // cell + 2
        g = rng.random((gh, gw))
        yy = np.arange(H) / cell
        xx = np.arange(W) / cell
        y0 = yy.astype(int); x0 = xx.astype(int)
        fy = (yy - y0)[:, None]; fx = (xx - x0)[None, :]
        a = g[y0][:, x0]; b = g[y0][:, x0 + 1]; c = g[y0 + 1][:, x0]; d = g[y0 + 1][:, x0 + 1]
        acc += (a * (1 - fy) * (1 - fx) + b * (1 - fy) * fx + c * fy * (1 - fx) + d * fy * fx) / (2 ** o)
    acc += 0.08 * rng.random((H, W)) * acc.max()
    acc -= acc.min()
    acc *= 255.0 / acc.max()
    return acc.astype(np.uint8)
