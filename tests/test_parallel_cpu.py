"""CPU, world size 2 over gloo: the multi-GPU host logic (range-slice sharding of one image, image-parallel batches,
gather of the winner records) with the CPU oracle standing in for the per-rank encode."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np
import pytest

from conftest import PKG, ROOT


def test_range_slices_partition_everything():
    from fic_b200.parallel import image_shares, range_slice
    for total in (0, 1, 7, 4096, 262144):
        for world in (1, 2, 3, 8):
            cuts = [range_slice(total, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == total
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            sizes = [e - b for b, e in cuts]
            assert max(sizes) - min(sizes) <= 1
    assert image_shares(5, 1, 2) == [3, 4] and image_shares(1, 1, 2) == []
    with pytest.raises(ValueError):
        range_slice(10, 2, 2)


WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, {pkg!r}); sys.path.insert(0, {root!r})
    import numpy as np, torch.distributed as dist
    from fic_b200.parallel import encode_sharded, encode_batch, decode_batch
    from oracle import oracle as orc
    dist.init_process_group("gloo")
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (48, 40), dtype=np.uint8)
    full = orc.bf_encode(img, mode=1, nthreads=1)
    got = encode_sharded(img, lambda im, b, e: orc.bf_encode(im, mode=1, range_begin=b, range_end=e, nthreads=1))
    assert all(np.array_equal(g, w) for g, w in zip(got, full)), "sharded != single"
    imgs = rng.integers(0, 256, (3, 32, 32), dtype=np.uint8)
    want = [orc.bf_encode(im, mode=1, nthreads=1) for im in imgs]
    idx, so, R = encode_batch(imgs, lambda im: orc.bf_encode(im, mode=1, nthreads=1))
    for b in range(3):
        assert np.array_equal(idx[b], want[b][0]) and np.array_equal(so[b], want[b][1]) and np.array_equal(R[b], want[b][2])
    dec = decode_batch(idx, so, 32, 32, lambda i, s: np.stack([orc.bf_decode(a, b_, 32, 32, 8) for a, b_ in zip(i, s)]))
    for b in range(3):
        assert np.array_equal(dec[b], orc.bf_decode(want[b][0], want[b][1], 32, 32, 8)), "batch decode != single"
    dist.destroy_process_group()
    print("rank", dist.get_rank() if dist.is_initialized() else os.environ["RANK"], "ok")
''')


def test_two_rank_gloo_sharding_matches_single_process(tmp_path):
    pytest.importorskip("torch")
    script = tmp_path / "worker.py"
    script.write_text(WORKER.format(pkg=PKG, root=ROOT))
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="1")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=300)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o
