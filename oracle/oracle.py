"""ctypes front-end of the CPU oracle (oracle/fic_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of fic_oracle.c.  Only tests/,
__graft_entry__.smoke() and bench.py's CPU legs may import this module; the
product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libfic_oracle.so")


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "fic_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def _u8(img):
    a = np.ascontiguousarray(img, dtype=np.uint8)
    assert a.ndim == 2
    return a


def num_threads() -> int:
    return int(lib().orc_num_threads())


def bf_encode(img, range_size=8, domain_size=16, strides=(1, 1), mode=1, range_begin=0, range_end=None,
              nthreads=0):
    """mode 1 = JitedBruteForceEncoder semantics, 0 = BruteForceEncoder (NumPy) semantics.
    Returns (idx[n,4] int32, so[n,2] float64, R[n] float64) for ranges [range_begin, range_end)."""
    img = _u8(img)
    H, W = img.shape
    total = (H // range_size) * (W // range_size)
    if range_end is None:
        range_end = total
    n = range_end - range_begin
    idx = np.zeros((n, 4), np.int32)
    so = np.zeros((n, 2), np.float64)
    R = np.zeros(n, np.float64)
    rc = lib().orc_bf_encode(_p(img, C.c_uint8), H, W, range_size, domain_size, int(strides[0]), int(strides[1]),
                             mode, range_begin, range_end, _p(idx, C.c_int32), _p(so, C.c_double),
                             _p(R, C.c_double), nthreads)
    if rc != 0:
        raise ValueError(f"orc_bf_encode failed: {rc}")
    return idx, so, R


def bf_score_one(img, ri, rj, di, dj, orient, rot, range_size=8, mode=1):
    img = _u8(img)
    H, W = img.shape
    out = np.zeros(3, np.float64)
    rc = lib().orc_bf_score_one(_p(img, C.c_uint8), H, W, range_size, mode, ri, rj, di, dj, orient, rot,
                                _p(out, C.c_double))
    if rc != 0:
        raise ValueError("orc_bf_score_one: bad arguments")
    return out  # R, s, o


def bf_decode(idx, so, width, height, iterations=32):
    idx = np.ascontiguousarray(idx, np.int32)
    so = np.ascontiguousarray(so, np.float64)
    out = np.zeros((height, width), np.uint8)
    rc = lib().orc_bf_decode(_p(idx, C.c_int32), _p(so, C.c_double), idx.shape[0], width, height, iterations,
                             _p(out, C.c_uint8))
    if rc != 0:
        raise ValueError("orc_bf_decode failed")
    return out


def qt_classify(img, si, sj, side, neg=False):
    img = _u8(img)
    out = np.zeros(4, np.int32)
    rc = lib().orc_qt_classify(_p(img, C.c_uint8), img.shape[0], img.shape[1], si, sj, side, int(neg),
                               _p(out, C.c_int32))
    if rc != 0:
        raise ValueError("orc_qt_classify: bad arguments")
    return tuple(int(v) for v in out)  # mc, vc, rot, orient


def qt_encode(img, scale_bits, offset_bits, max_scale, full_class_search, only_positive_scale, min_range_side,
              max_range_side, error_tolerance, domain_hstep=-1, domain_vstep=-1):
    """Returns (nodes[k,8] int32, vals[k,3] float64, n_roots): preorder stream, see orc_qt_encode."""
    img = _u8(img)
    H, W = img.shape
    cap = 2 * H * W + 64
    nodes = np.zeros((cap, 8), np.int32)
    vals = np.zeros((cap, 3), np.float64)
    nroots = C.c_int32(0)
    f = lib().orc_qt_encode
    f.argtypes = [C.POINTER(C.c_uint8), C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int,
                  C.c_int, C.c_double, C.c_int, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_double), C.c_int,
                  C.POINTER(C.c_int32)]
    rc = f(_p(img, C.c_uint8), H, W, scale_bits, offset_bits, float(max_scale), int(full_class_search),
           int(only_positive_scale), min_range_side, max_range_side, float(error_tolerance), domain_hstep,
           domain_vstep, _p(nodes, C.c_int32), _p(vals, C.c_double), cap, C.byref(nroots))
    if rc < 0:
        raise ValueError(f"orc_qt_encode failed: {rc}")
    return nodes[:rc].copy(), vals[:rc].copy(), int(nroots.value)


def hv_encode(img, min_range_height, min_range_width, max_range_height, max_range_width, scale_bits, offset_bits,
              err_threshold, max_scale=1.0, nthreads=0):
    img = _u8(img)
    H, W = img.shape
    cap = 2 * H * W + 64
    nodes = np.zeros((cap, 8), np.int32)
    vals = np.zeros((cap, 3), np.float64)
    f = lib().orc_hv_encode
    f.argtypes = [C.POINTER(C.c_uint8), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                  C.c_double, C.c_double, C.POINTER(C.c_int32), C.POINTER(C.c_double), C.c_int, C.c_int]
    rc = f(_p(img, C.c_uint8), H, W, min_range_height, min_range_width, max_range_height, max_range_width,
           scale_bits, offset_bits, float(err_threshold), float(max_scale), _p(nodes, C.c_int32),
           _p(vals, C.c_double), cap, nthreads)
    if rc < 0:
        raise ValueError(f"orc_hv_encode failed: {rc}")
    return nodes[:rc].copy(), vals[:rc].copy()


def hv_search(img, i, j, w, h, scale_bits, offset_bits, max_scale=1.0):
    img = _u8(img)
    rec = np.zeros(6, np.int32)
    vals = np.zeros(3, np.float64)
    f = lib().orc_hv_search
    f.argtypes = [C.POINTER(C.c_uint8), C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                  C.c_double, C.POINTER(C.c_int32), C.POINTER(C.c_double)]
    f(_p(img, C.c_uint8), img.shape[0], img.shape[1], i, j, w, h, scale_bits, offset_bits, float(max_scale),
      _p(rec, C.c_int32), _p(vals, C.c_double))
    return rec, vals  # (di,dj,orient,rot,qs,qo), (s,o,err)


def tree_decode(kind, leaves, so, H, W, iterations=-1, stop_rel_err=5e-3, min_it=2, max_it=10, f32_affine=False,
                deblock=False):
    """kind 0 = quadtree, 1 = HV.  leaves[k] = (ri,rj,h,w,di,dj,orient,rot); so[k] = (scale, offset).
    deblock: quadtree only, leaves in the reference's visiting order (see qt_visit_order)."""
    leaves = np.ascontiguousarray(leaves, np.int32)
    so = np.ascontiguousarray(so, np.float64)
    out = np.zeros((H, W), np.uint8)
    done = C.c_int32(0)
    f = lib().orc_tree_decode_post
    f.argtypes = [C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_double), C.c_int, C.c_int, C.c_int, C.c_int,
                  C.c_double, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8), C.POINTER(C.c_int32)]
    rc = f(kind, _p(leaves, C.c_int32), _p(so, C.c_double), leaves.shape[0], H, W, iterations, float(stop_rel_err),
           min_it, max_it, int(f32_affine), int(bool(deblock)), _p(out, C.c_uint8), C.byref(done))
    if rc != 0:
        raise ValueError("orc_tree_decode failed")
    return out, int(done.value)


def qt_visit_order(nodes, n_roots, H, W):
    """Leaf rectangles (ri, rj, size) of a preorder quadtree stream in the order the reference's decoder and
    post-processor visit them (quadtree/postprocessing.py:29-71): roots in _de_partition order, each root breadth
    first.  Returns (rects[k,3], stream_index[k])."""
    from collections import deque
    kinds = np.asarray(nodes)[:, 0].tolist()
    roots = []

    def part(i, j, w, h):
        e = min(w.bit_length(), h.bit_length()) - 1
        side = 1 << e
        roots.append((i, j, side))
        if side < w:
            part(i, j + side, w - side, h)
        if side < h:
            part(i + side, j, side, h - side)

    part(0, 0, W, H)
    assert len(roots) == n_roots
    # subtree extents in the stream, so a child's stream index can be found without rebuilding objects
    n = len(kinds)
    end = [0] * n
    for k in range(n - 1, -1, -1):
        if kinds[k]:
            end[k] = k + 1
        else:
            c = k + 1
            for _ in range(4):
                c = end[c]
            end[k] = c
    rects, index = [], []
    pos = 0
    for (ri, rj, side) in roots:
        queue = deque([(pos, ri, rj, side)])
        pos = end[pos]
        while queue:
            k, i, j, size = queue.popleft()
            if kinds[k]:
                rects.append((i, j, size))
                index.append(k)
            else:
                half = size // 2
                c = k + 1
                for (di, dj) in ((0, 0), (0, half), (half, 0), (half, half)):
                    queue.append((c, i + di, j + dj, half))
                    c = end[c]
    return np.array(rects, np.int32).reshape(-1, 3), np.array(index, np.int64)


def qt_deblock(img, rects, weights=(5. / 6., 1. - 5. / 6., 2. / 3., 1. - 2. / 3.), small_side=4):
    """DeblockingQuadtreePostprocessor.postprocess on a float64 image; rects[k] = (ri, rj, size) in visiting order."""
    a = np.ascontiguousarray(img, np.float64)
    r = np.asarray(rects, np.int32).reshape(-1, 3)
    leaves = np.zeros((r.shape[0], 8), np.int32)
    leaves[:, 0], leaves[:, 1], leaves[:, 2], leaves[:, 3] = r[:, 0], r[:, 1], r[:, 2], r[:, 2]
    w = np.array(weights, np.float64)
    out = np.empty_like(a)
    f = lib().orc_qt_deblock
    f.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_int32), C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double), C.c_int,
                  C.POINTER(C.c_double)]
    f(_p(a, C.c_double), _p(leaves, C.c_int32), leaves.shape[0], a.shape[0], a.shape[1], _p(w, C.c_double),
      int(small_side), _p(out, C.c_double))
    return out


def rgb_to_ycbcr(rgb, subsample=False):
    a = np.ascontiguousarray(rgb, np.uint8)
    N, M, _ = a.shape
    cs = (N // 2, M // 2 + M % 2) if subsample else (N, M)
    y, cb, cr = np.zeros((N, M), np.uint8), np.zeros(cs, np.uint8), np.zeros(cs, np.uint8)
    f = lib().orc_rgb_to_ycbcr
    f.argtypes = [C.POINTER(C.c_uint8), C.c_int, C.c_int, C.c_int] + [C.POINTER(C.c_uint8)] * 3
    if f(_p(a, C.c_uint8), N, M, int(bool(subsample)), _p(y, C.c_uint8), _p(cb, C.c_uint8), _p(cr, C.c_uint8)) != 0:
        raise ValueError("orc_rgb_to_ycbcr: 4:2:0 needs an even height")
    return y, cb, cr


def ycbcr_to_rgb(y, cb, cr, subsample=False):
    y, cb, cr = (np.ascontiguousarray(p, np.uint8) for p in (y, cb, cr))
    N, M = y.shape
    out = np.zeros((N, M, 3), np.uint8)
    f = lib().orc_ycbcr_to_rgb
    f.argtypes = [C.POINTER(C.c_uint8)] * 3 + [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_uint8)]
    if f(_p(y, C.c_uint8), _p(cb, C.c_uint8), _p(cr, C.c_uint8), N, M, int(bool(subsample)), _p(out, C.c_uint8)) != 0:
        raise ValueError("orc_ycbcr_to_rgb: 4:2:0 needs a height that is a multiple of 4")
    return out


def average_subsample(arr, jit=False):
    a = np.ascontiguousarray(arr, np.float64)
    out = np.zeros((a.shape[0] // 2, a.shape[1] // 2), np.float64)
    rc = lib().orc_average_subsample(_p(a, C.c_double), a.shape[0], a.shape[1], int(jit), _p(out, C.c_double))
    if rc != 0:
        raise ValueError("orc_average_subsample: odd shape")
    return out
