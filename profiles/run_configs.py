#!/usr/bin/env python
"""Times the five BASELINE.json configs on one B200 and checks each against the exact path / the CPU oracle.
Run on the GPU box:  python profiles/run_configs.py > gpurun_out/configs.jsonl
(Not a bench.py replacement: bench.py holds the headline contract; this records the other configs for the notes.)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "fractal-image-compression_b200"), ROOT, os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import fic_b200  # noqa: E402
from conftest import synthetic_image  # noqa: E402
from oracle import oracle as orc  # noqa: E402

imgs = np.load(os.path.join(ROOT, "tests/golden/images.npz"))
h = fic_b200.default_handle(0)


def timed(fn, reps=3):
    fn()
    best = 1e30
    for _ in range(reps):
        t = time.perf_counter(); out = fn(); best = min(best, time.perf_counter() - t)
    return best, out


def psnr(a, b):
    return float(10 * np.log10(255.0 ** 2 / np.mean((a.astype(np.float64) - b.astype(np.float64)) ** 2)))


def emit(**kw):
    print(json.dumps(kw), flush=True)


# config 1: 256x256 brute force encode + decode
img = imgs["lena256"]
t_enc, (idx, so, R) = timed(lambda: fic_b200.bf_encode(img, mode=1))
t_dec, dec = timed(lambda: fic_b200.bf_decode(idx, so, 256, 256, 32))
from bruteforce.decoder import Decoder  # noqa: E402
from bruteforce.jit_bfencoder import JitedBruteForceEncoder  # noqa: E402
t_cls_enc, (enc_obj, errs) = timed(lambda: JitedBruteForceEncoder(img, 8, 16).encode())
t_cls_dec, dec_cls = timed(lambda: Decoder().decode(enc_obj, 32))
emit(config=1, what="bruteforce-jit 256x256 encode+decode (host buffers)", encode_ms=t_enc * 1e3, decode_ms=t_dec * 1e3,
     class_api_encode_ms=t_cls_enc * 1e3, class_api_decode_ms=t_cls_dec * 1e3, class_api_equals_c_abi=bool(np.array_equal(dec_cls, dec)),
     psnr_db=psnr(dec, img), matches_per_s=1024 * 240 * 240 * 8 / t_enc, reference_cpu_s="415.5 (1 core, BASELINE.md)")

# config 2: quadtree 512x512
from quadtree.common import forest_to_leaves  # noqa: E402
from quadtree.decoder import QuadtreeDecoder  # noqa: E402
from quadtree.encoder import QuadtreeEncoder  # noqa: E402
img = imgs["lena512"]
qp = (7, 5, 1.0, False, False, 4, 8, 2.0)
t_gpu, q = timed(lambda: QuadtreeEncoder(*qp).encode(img))
t_raw, _ = timed(lambda: fic_b200.qt_encode(img, *qp))
t_cpu, ref = timed(lambda: orc.qt_encode(img, *qp), reps=1)
t_qdec, qdec = timed(lambda: QuadtreeDecoder().decode(q))
from quadtree.serialization import QuadtreeDeserializer, QuadtreeSerializer  # noqa: E402
import io  # noqa: E402
from entropy.processor import CodingDimension, HuffmanEntropyProcessor  # noqa: E402


def qt_to_file():
    out = io.BytesIO()
    QuadtreeSerializer(HuffmanEntropyProcessor(), set(CodingDimension)).serialize(QuadtreeEncoder(*qp).encode(img), out)
    return out.getvalue()


t_file, qblob = timed(qt_to_file)
t_back, qback = timed(lambda: QuadtreeDecoder().decode(QuadtreeDeserializer().deserialize(io.BytesIO(qblob))))
emit(config=2, what="quadtree 512x512 (7,5,1.0,False,False,4,8,2.0)", encode_ms_class_api=t_gpu * 1e3, encode_ms_c_abi=t_raw * 1e3,
     image_to_fqt_bytes_ms=t_file * 1e3, fqt_bytes=len(qblob), fqt_bytes_to_image_ms=t_back * 1e3,
     file_round_trip_equals_decode=bool(np.array_equal(qback, qdec)),
     oracle_cpu_ms=t_cpu * 1e3, decode_ms=t_qdec * 1e3, psnr_db=psnr(qdec, img), nodes=int(len(ref[0])),
     reference_cpu_s="3.74 (BASELINE.md)")

# config 2 stress rows: full class search, and dense domain pools (step 1)
for tag, params in (("full_class_search", (7, 5, 1.0, True, False, 4, 8, 2.0)),
                    ("dense_pools_step1", (7, 5, 1.0, False, False, 4, 8, 2.0, 1, 1)),
                    ("dense_pools_step1_full_class_search", (7, 5, 1.0, True, False, 4, 8, 2.0, 1, 1))):
    t_s, qs = timed(lambda: QuadtreeEncoder(*params).encode(img), reps=2)
    t_d, ds = timed(lambda: QuadtreeDecoder().decode(qs), reps=2)
    emit(config="2-stress", what=f"quadtree 512x512 {tag} {params}", encode_ms=t_s * 1e3, decode_ms=t_d * 1e3,
         psnr_db=psnr(ds, img), leaves=int(forest_to_leaves(qs)[0].shape[0]))

# config 3: HV 512x512
from hvv2.decoder import HVDecoder  # noqa: E402
from hvv2.encoder import HVEncoder  # noqa: E402
hp = (4, 4, 16, 16, 5, 7, 25.0)
t_gpu, hv = timed(lambda: HVEncoder(*hp).encode(img), reps=2)
t_hdec, hdec = timed(lambda: HVDecoder().decode(hv))
g = np.load(os.path.join(ROOT, "tests/golden/hv_golden.npz"))
leaves = int((g["lena512_nodes"][:, 0] == 1).sum()) if "lena512_nodes" in g.files else None
import io  # noqa: E402
from entropy.processor import CodingDimension, HuffmanEntropyProcessor  # noqa: E402
from hvv2.serialization import HVSerializer  # noqa: E402


def hv_to_bytes():
    out = io.BytesIO()
    HVSerializer(HuffmanEntropyProcessor(), set(CodingDimension)).serialize(hv, out)
    return out.getvalue()


t_ser, blob = timed(hv_to_bytes, reps=3)
from hvv2.serialization import HVDeserializer  # noqa: E402
t_hback, hback = timed(lambda: HVDecoder().decode(HVDeserializer().deserialize(io.BytesIO(blob))))
emit(config=3, what="HV 512x512 (4,4,16,16,5,7,25.0)", encode_ms=t_gpu * 1e3, entropy_coded_stream_ms=t_ser * 1e3,
     stream_bytes=len(blob), decode_ms=t_hdec * 1e3, file_bytes_to_image_ms=t_hback * 1e3,
     file_round_trip_equals_decode=bool(np.array_equal(hback, hdec)), psnr_db=psnr(hdec, img),
     leaves_reference=leaves, reference_cpu_s="1688 (BASELINE.md), 1578 when the golden file was made")

# config 4: synthetic 4096x4096 brute force on one GPU (+ a 1/8 range slice as one rank of eight would run it)
big = synthetic_image(0, 4096, 4096)
n = 512 * 512
h.reset_counters()
t0 = time.perf_counter(); idx, so, R = fic_b200.bf_encode(big, mode=1); t_full = time.perf_counter() - t0
c = h.counters()
t0 = time.perf_counter(); part = fic_b200.bf_encode(big, mode=1, range_begin=3 * n // 8, range_end=4 * n // 8); t_slice = time.perf_counter() - t0
ok = np.array_equal(part[0], idx[3 * n // 8:4 * n // 8]) and np.array_equal(part[1], so[3 * n // 8:4 * n // 8])
sel = [0, 777, 131071, 200000, n - 1]
exact_ok = True
for k in sel:
    w = fic_b200.bf_encode(big, mode=1, flags=fic_b200.BF_EXACT_ONLY, range_begin=k, range_end=k + 1)
    exact_ok &= bool(np.array_equal(w[0][0], idx[k]) and w[1][0, 0] == so[k, 0] and w[1][0, 1] == so[k, 1] and w[2][0] == R[k])
matches = n * 4080 * 4080 * 8
emit(config=4, what="bruteforce-jit synthetic 4096x4096, all 262144 ranges on ONE GPU", seconds=t_full, matches_per_s=matches / t_full,
     slice_1_of_8_seconds=t_slice, slice_equals_full=bool(ok), five_ranges_equal_exact_scan=exact_ok, counters=c)

# config 5: one 1024x1024 RGB image, per channel brute force (batch of 256 = 256x this, image-parallel)
rgb = np.stack([synthetic_image(100 + k, 1024, 1024) for k in range(3)])
t_img, out = timed(lambda: fic_b200.bf_encode_batch(rgb, mode=1), reps=2)
t_dec3, dec3 = timed(lambda: fic_b200.bf_decode_batch(out[0], out[1], 1024, 1024, 32), reps=1)
# a rank's share of the 256-image batch on 8 GPUs: 32 images = 96 planes, decoded by one batched call
share_idx = np.concatenate([out[0]] * 32); share_so = np.concatenate([out[1]] * 32)
t_dec96, dec96 = timed(lambda: fic_b200.bf_decode_batch(share_idx, share_so, 1024, 1024, 32), reps=1)
emit(config=5, what="one synthetic 1024x1024x3 image, 3 planes brute force", encode_s=t_img, decode_s=t_dec3,
     decode_96_planes_s=t_dec96, planes_equal=bool(np.array_equal(dec96[3:6], dec3)),
     matches_per_s=3 * 16384 * 1008 * 1008 * 8 / t_img, batch_of_256_on_8_gpus_estimate_s=32 * t_img + t_dec96)

# config 5 (ii): the reference's own colour entry point (quadtree/rgb.py:20-27) with the config-2 parameters, planes kept
# in device memory; decode plain and with the stock deblocking filter fused into the decoder
from quadtree.common import forest_to_leaves  # noqa: E402
from quadtree.postprocessing import DeblockingQuadtreePostprocessor  # noqa: E402
from quadtree.rgb import RGBQuadtreeDecoder, RGBQuadtreeEncoder  # noqa: E402
rgb_hwc = np.ascontiguousarray(rgb.transpose(1, 2, 0))
for mode in ("4:4:4", "4:2:0"):
    enc = RGBQuadtreeEncoder(QuadtreeEncoder(*qp), QuadtreeEncoder(*qp))
    t_enc, col = timed(lambda: enc.encode(rgb_hwc, mode), reps=2)
    dec = RGBQuadtreeDecoder(QuadtreeDecoder())
    t_dec, back = timed(lambda: dec.decode(col), reps=2)
    t_decdb, backdb = timed(lambda: dec.decode(col, DeblockingQuadtreePostprocessor()), reps=2)
    t_conv, planes = timed(lambda: fic_b200.rgb_to_ycbcr(rgb_hwc, mode), reps=3)
    emit(config="5(ii)", what=f"RGBQuadtreeEncoder(7,5,1.0,..,4,8,2.0) 1024x1024x3 {mode}", encode_ms=t_enc * 1e3,
         decode_ms=t_dec * 1e3, decode_deblocked_ms=t_decdb * 1e3, colour_conversion_host_arrays_ms=t_conv * 1e3,
         leaves=[int(forest_to_leaves(p)[0].shape[0]) for p in (col.encoded_Y, col.encoded_Cb, col.encoded_Cr)],
         psnr_db=psnr(back.astype(np.float64), rgb_hwc), psnr_deblocked_db=psnr(backdb.astype(np.float64), rgb_hwc))
