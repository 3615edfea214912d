#!/usr/bin/env python
"""Generate the golden fixtures in this directory by running the UNMODIFIED reference.

Run in the build container only (needs /root/reference, numba, Pillow):

    python tests/golden/make_golden.py [--slow]

Outputs (committed):
  images.npz          grayscale test inputs (load_grayscale of imgs/raw/*), uint8
  bf_golden.npz       brute-force encoders/decoder on crops (+ full 256x256 with --slow)
  qt_golden.npz       quadtree encoder node streams + decoded images
  hv_golden.npz       HV encoder node streams + decoded images (+ 256/512 with --slow)
  *.fqt               the reference's bundled compressed files (golden byte streams)

Node streams are preorder: quadtree node = [kind, di, dj, orient, rot, qs, qo], kind 1 = leaf,
0 = internal (followed by its 4 children); HV internal = [0, vertical, split_idx, 0, 0, 0, 0].
The GPU box has no /root/reference; tests read only these files.
"""
import argparse
import hashlib
import os
import shutil
import struct
import sys
import time

import numpy as np

REF = "/root/reference"
sys.path.insert(0, os.path.join(REF, "src"))
HERE = os.path.dirname(os.path.abspath(__file__))

from utils import load_grayscale  # noqa: E402  (reference module)


def raw(name):
    return load_grayscale(os.path.join(REF, "imgs/raw", name))


def qt_flatten(qimg):
    nodes, vals = [], []

    def rec(n):
        if n.is_leaf():
            d = n.domain
            nodes.append([1, d.start_i, d.start_j, d.orientation, d.rotation, d.quantized_scale, d.quantized_offset])
            vals.append([d.scale, d.offset])
        else:
            nodes.append([0] * 7)
            vals.append([0.0, 0.0])
            for c in n.children:
                rec(c)

    for r in qimg.forest:
        rec(r)
    return np.array(nodes, np.int32), np.array(vals, np.float64), len(qimg.forest)


def hv_flatten(h):
    nodes, vals = [], []
    stack = [h.root]
    while stack:  # iterative preorder (trees are up to ~70 deep, recursion would also do)
        n = stack.pop()
        if n.is_leaf():
            d = n.domain
            nodes.append([1, d.start_i, d.start_j, d.orientation, d.rotation, d.quantized_scale, d.quantized_offset])
            vals.append([d.scale, d.offset])
        else:
            nodes.append([0, int(n.split_info.vertical_split), int(n.split_info.split_idx), 0, 0, 0, 0])
            vals.append([0.0, 0.0])
            stack.append(n.c2)
            stack.append(n.c1)
    return np.array(nodes, np.int32), np.array(vals, np.float64)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--slow", action="store_true", help="also run the multi-minute reference encodes")
    args = ap.parse_args()

    imgs = {
        "lena256": raw("lena_256x256_grayscale.bmp"),
        "cauliflower256": raw("cauliflower_256x256_grayscale.png"),
        "forest256": raw("forest_256x256_grayscale.png"),
        "lena512": raw("lena_512x512_colored.bmp"),
        "lena313x413": raw("lena_313x413_colored.bmp"),
        "cauliflower64": raw("cauliflower_64x64_colored.png"),
        "cauliflower128": raw("cauliflower_128x128_colored.png"),
        "cauliflower240x92": raw("cauliflower_240x92_colored.png"),
    }
    np.savez_compressed(os.path.join(HERE, "images.npz"), **imgs)
    for f in ("kalafior256.fqt", "lena313x413.fqt"):
        shutil.copyfile(os.path.join(REF, "imgs/compressed", f), os.path.join(HERE, f))

    # ---------------- brute force ----------------
    from bruteforce.bf_encoder import BruteForceEncoder
    from bruteforce.decoder import Decoder
    from bruteforce.jit_bfencoder import JitedBruteForceEncoder

    out = {}
    lena = imgs["lena256"]

    def bf_pack(enc):
        idx = np.array([[d.start_i, d.start_j, d.orientation, d.rotation] for d in enc.domains], np.int32)
        so = np.array([[d.s, d.o] for d in enc.domains], np.float64)
        return idx, so

    crop48 = np.ascontiguousarray(lena[100:148, 60:108])
    enc, errs = JitedBruteForceEncoder(crop48).encode()
    out["jit48_img"] = crop48
    out["jit48_idx"], out["jit48_so"] = bf_pack(enc)
    out["jit48_errs"] = np.array(errs, np.float64)
    out["jit48_dec32"] = Decoder().decode(enc, 32)
    out["jit48_dec3"] = Decoder().decode(enc, 3)

    # flat / tie-heavy synthetic: tie-break must follow scan order
    rng = np.random.default_rng(7)
    flat = np.full((48, 48), 90, np.uint8)
    flat[8:24, 8:40] = 200
    flat[30:40, 4:20] = rng.integers(0, 255, (10, 16))
    enc, errs = JitedBruteForceEncoder(flat).encode()
    out["jitflat_img"] = flat
    out["jitflat_idx"], out["jitflat_so"] = bf_pack(enc)
    out["jitflat_errs"] = np.array(errs, np.float64)
    out["jitflat_dec32"] = Decoder().decode(enc, 32)

    # extreme values (all-255 rows / checkerboard): exactness of the integer contraction
    chk = (np.indices((48, 48)).sum(0) % 2 * 255).astype(np.uint8)
    chk[:16] = 255
    chk[40:, 10:30] = rng.integers(0, 255, (8, 20))
    enc, errs = JitedBruteForceEncoder(chk).encode()
    out["jitchk_img"] = chk
    out["jitchk_idx"], out["jitchk_so"] = bf_pack(enc)
    out["jitchk_errs"] = np.array(errs, np.float64)

    crop32 = np.ascontiguousarray(lena[100:132, 60:92])
    enc = BruteForceEncoder(crop32).encode()
    out["np32_img"] = crop32
    out["np32_idx"], out["np32_so"] = bf_pack(enc)
    out["np32_dec32"] = Decoder().decode(enc, 32)
    enc = BruteForceEncoder(crop32, strides=(2, 3)).encode()
    out["np32s23_idx"], out["np32s23_so"] = bf_pack(enc)

    # 512x512 headline image: a few ranges only (1.7 s per range in the reference)
    from bruteforce.jit_bfencoder import _find_best_domain

    l512 = imgs["lena512"]
    sel = np.array([0, 37, 1000, 2079, 2080, 3333, 4095], np.int32)
    rec = []
    for k in sel:
        i, j = (k // 64) * 8, (k % 64) * 8
        rec.append(_find_best_domain(l512, l512[i:i + 8, j:j + 8], 16))
    out["jit512_sel"] = sel
    out["jit512_rec"] = np.array(rec, np.float64)  # di,dj,orient,rot,s,o,R

    if args.slow:
        t = time.time()
        enc, errs = JitedBruteForceEncoder(lena).encode()
        print("full 256 jit encode", time.time() - t, "s")
        idx, so = bf_pack(enc)
        out["jit256_idx"], out["jit256_so"] = idx, so
        out["jit256_errs"] = np.array(errs, np.float64)
        dec = Decoder().decode(enc, 32)
        out["jit256_dec32"] = dec
        h = hashlib.sha256()
        for (a, b, c, d), (s, o) in zip(idx.tolist(), so.tolist()):
            h.update(struct.pack("<iiiidd", a, b, c, d, s, o))
        print("records sha256", h.hexdigest())  # a6f53f58...
        print("decoded sha256", hashlib.sha256(dec.tobytes()).hexdigest())
    else:
        old = os.path.join(HERE, "bf_golden.npz")
        if os.path.exists(old):
            prev = np.load(old)
            for k in prev.files:
                if k.startswith("jit256_"):
                    out[k] = prev[k]
    np.savez_compressed(os.path.join(HERE, "bf_golden.npz"), **out)

    # ---------------- quadtree ----------------
    from quadtree.decoder import QuadtreeDecoder
    from quadtree.encoder import QuadtreeEncoder

    out = {}
    qt_cases = {
        "cauliflower256": (7, 5, 1.0, False, False, 4, 8, 2.0),       # == kalafior256.fqt
        "lena313x413": (7, 5, 1.0, False, False, 4, 8, 2.0),          # == lena313x413.fqt
        "lena512": (7, 5, 1.0, False, False, 4, 8, 2.0),              # config 2
        "cauliflower64": (7, 5, 1.0, True, False, 2, 16, 4.0),        # full class search
        "cauliflower128": (5, 7, 0.9, False, True, 2, 32, 6.0),       # positive only, max_scale 0.9
        "cauliflower240x92": (7, 5, 1.0, False, False, 4, 16, 3.0),   # non-square forest
    }
    names = []
    for name, params in qt_cases.items():
        q = QuadtreeEncoder(*params).encode(imgs[name])
        nodes, vals, nroots = qt_flatten(q)
        out[f"{name}_params"] = np.array(params, np.float64)
        out[f"{name}_nodes"], out[f"{name}_vals"] = nodes, vals
        out[f"{name}_nroots"] = np.array(nroots)
        out[f"{name}_dec"] = QuadtreeDecoder().decode(q)
        out[f"{name}_dec4"] = QuadtreeDecoder(iterations=4).decode(q)
        names.append(name)
    # dense domain pool (hstep=vstep=4) on a small image
    q = QuadtreeEncoder(7, 5, 1.0, False, False, 4, 8, 2.0, 4, 4).encode(imgs["cauliflower64"])
    nodes, vals, nroots = qt_flatten(q)
    out["cauliflower64step4_params"] = np.array((7, 5, 1.0, 0, 0, 4, 8, 2.0, 4, 4), np.float64)
    out["cauliflower64step4_nodes"], out["cauliflower64step4_vals"] = nodes, vals
    out["cauliflower64step4_nroots"] = np.array(nroots)
    out["cauliflower64step4_dec"] = QuadtreeDecoder().decode(q)
    np.savez_compressed(os.path.join(HERE, "qt_golden.npz"), **out)

    # ---------------- HV ----------------
    from hvv2.decoder import HVDecoder
    from hvv2.encoder import HVEncoder

    out = {}
    hv_cases = {
        "lena64": (np.ascontiguousarray(lena[64:128, 64:128]), (4, 4, 16, 16, 5, 7, 25.0)),
        "lena48x80": (np.ascontiguousarray(lena[100:148, 60:140]), (2, 2, 32, 32, 5, 7, 16.0)),
        "cauliflower64": (imgs["cauliflower64"], (4, 4, 16, 16, 5, 7, 25.0, 0.8)),
        "cauliflower240x92": (imgs["cauliflower240x92"], (3, 5, 24, 40, 6, 6, 30.0)),
    }
    if args.slow:
        hv_cases["lena256"] = (lena, (4, 4, 16, 16, 5, 7, 25.0))
        hv_cases["lena512"] = (imgs["lena512"], (4, 4, 16, 16, 5, 7, 25.0))   # config 3, ~28 min
    for name, (img, params) in hv_cases.items():
        t = time.time()
        h = HVEncoder(*params).encode(img)
        print("hv", name, time.time() - t, "s")
        nodes, vals = hv_flatten(h)
        out[f"{name}_img"] = img
        out[f"{name}_params"] = np.array(params, np.float64)
        out[f"{name}_nodes"], out[f"{name}_vals"] = nodes, vals
        out[f"{name}_dec"] = HVDecoder().decode(h)
        out[f"{name}_dec5nq"] = HVDecoder(iterations=5).using_not_quantized_values().decode(h)
    if not args.slow:
        old = os.path.join(HERE, "hv_golden.npz")
        if os.path.exists(old):
            prev = np.load(old)
            for k in prev.files:
                if k.startswith("lena256_") or k.startswith("lena512_"):
                    out[k] = prev[k]
    np.savez_compressed(os.path.join(HERE, "hv_golden.npz"), **out)
    print("done")


if __name__ == "__main__":
    main()
