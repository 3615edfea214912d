"""ctypes binding of libfic_b200.so -- the sm_100a implementation of the block-matching hot path.

This is the only module that touches the shared library.  The reference-facing classes
(``bruteforce``, ``quadtree``, ``hvv2`` next to this package) call the thin functions below with
NumPy arrays; ``bench.py`` additionally uses the ``*_dev`` forms with raw device pointers
(``torch.Tensor.data_ptr()``).

There is deliberately no CPU fallback: if the library is missing, or no sm_100 GPU is present,
the first call raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfic_b200.so")

DEVICE_PTRS = 1
BF_EXACT_ONLY = 2
BF_NO_FALLBACK = 4
QT_DEBLOCK = 8
TREE_OUT_F64 = 16
BF_MODE_NUMPY = 0
BF_MODE_JIT = 1
TREE_QUADTREE = 0
TREE_HV = 1
ITER_AUTO = -1

_u8p = C.POINTER(C.c_uint8)
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)
_vp = C.c_void_p

# name -> (restype, argtypes); must list every symbol of include/fic.h (tests check this)
SIGNATURES = {
    "fic_version": (C.c_char_p, []),
    "fic_last_error": (C.c_char_p, []),
    "fic_device_count": (C.c_int, []),
    "fic_create": (C.c_int, [C.c_int, C.POINTER(_vp)]),
    "fic_destroy": (C.c_int, [_vp]),
    "fic_get_counters": (C.c_int, [_vp, _i64p]),
    "fic_reset_counters": (C.c_int, [_vp]),
    "fic_synchronize": (C.c_int, [_vp]),
    "fic_profile": (C.c_int, [_vp, C.c_int]),
    "fic_profile_read": (C.c_int, [_vp, _f64p, _i64p]),
    "fic_bf_encode": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_int, _vp, _vp, _vp, C.c_uint, _vp]),
    "fic_bf_encode_batch": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp,
                                      C.c_uint, _vp]),
    "fic_bf_encode_static": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp,
                                       _vp, _vp]),
    "fic_bf_screen_probe": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.c_uint, _vp]),
    "fic_bf_decode": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.c_uint, _vp]),
    "fic_bf_decode_batch": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.c_uint, _vp]),
    "fic_average_subsample": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_uint, _vp]),
    "fic_qt_encode": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int,
                                C.c_int, C.c_double, C.c_int, C.c_int, _vp, _vp, C.c_int, _i32p, C.c_uint, _vp]),
    "fic_qt_classify": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _vp, C.c_int, _vp, C.c_uint, _vp]),
    "fic_hv_encode": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.c_double, C.c_double, _vp, _vp, C.c_int, C.c_uint, _vp]),
    "fic_hv_search": (C.c_int, [_vp, _vp, C.c_int, C.c_int, _vp, C.c_int, C.c_int, C.c_int, C.c_double, _vp, _vp,
                                C.c_uint, _vp]),
    "fic_tree_decode": (C.c_int, [_vp, C.c_int, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int,
                                  C.c_int, C.c_int, _vp, _i32p, C.c_uint, _vp]),
    "fic_qt_deblock": (C.c_int, [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_int, _vp, C.c_uint, _vp]),
    "fic_rgb_to_ycbcr": (C.c_int, [_vp, _vp, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, C.c_uint, _vp]),
    "fic_ycbcr_to_rgb": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, _vp, C.c_uint, _vp]),
    "fic_tree_read_nodes": (C.c_int, [C.c_int, _vp, C.c_int64, C.c_int64, _vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, C.c_int,
                                      _i64p]),
    "fic_hv_layout": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp, _vp]),
    "fic_huffman_read_tree": (C.c_int, [_vp, C.c_int64, C.c_int64, C.c_int, _vp, C.c_int, _i64p]),
    "fic_pack_fields": (C.c_int64, [_vp, _vp, C.c_int64, C.c_uint32, C.c_int, _vp, C.c_int64, C.POINTER(C.c_uint32),
                                    C.POINTER(C.c_int)]),
}

_lib = None
_lock = threading.Lock()


class FicError(RuntimeError):
    """CUDA / capacity failure inside libfic_b200.so."""


def lib():
    """Load the shared library (no GPU needed for loading, only for fic_create)."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise ImportError(
                    f"{LIB_PATH} is missing: build it with `python __graft_entry__.py` "
                    "(or `make -C fractal-image-compression_b200/csrc`). There is no CPU fallback.")
            L = C.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(L, name)
                fn.restype = res
                fn.argtypes = args
            _lib = L
    return _lib


def last_error() -> str:
    return lib().fic_last_error().decode("utf-8", "replace")


def _check(rc: int, what: str) -> int:
    if rc >= 0:
        return rc
    msg = f"{what}: {last_error()} (code {rc})"
    if rc == -1:
        raise ValueError(msg)      # where the reference raises / crashes on bad input
    if rc == -5:
        raise MemoryError(msg)
    raise FicError(msg)


class Handle:
    """One handle per GPU; used by one host thread at a time (include/fic.h)."""

    def __init__(self, device: int = 0):
        self._h = _vp()
        _check(lib().fic_create(int(device), C.byref(self._h)), "fic_create")
        self.device = int(device)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            lib().fic_destroy(self._h)
            self._h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def ptr(self):
        return self._h

    def counters(self):
        out = (C.c_int64 * 8)()
        _check(lib().fic_get_counters(self._h, out), "fic_get_counters")
        return {"launches": out[0], "screen_ranges": out[1], "rescan_ranges": out[2], "survivors": out[3],
                "h2d_bytes": out[4], "d2h_bytes": out[5]}

    def reset_counters(self):
        _check(lib().fic_reset_counters(self._h), "fic_reset_counters")

    def synchronize(self):
        _check(lib().fic_synchronize(self._h), "fic_synchronize")

    def profile(self, enable: bool):
        _check(lib().fic_profile(self._h, int(bool(enable))), "fic_profile")

    def profile_read(self):
        """(summed milliseconds, launches) of the dominant kernel since the last read."""
        ms = C.c_double(0.0)
        n = C.c_int64(0)
        _check(lib().fic_profile_read(self._h, C.byref(ms), C.byref(n)), "fic_profile_read")
        return float(ms.value), int(n.value)


_handles: dict[int, Handle] = {}


def default_handle(device: int | None = None) -> Handle:
    """Process-wide handle for `device` (default: $FIC_DEVICE, else $LOCAL_RANK, else 0)."""
    if device is None:
        device = int(os.environ.get("FIC_DEVICE", os.environ.get("LOCAL_RANK", "0")))
    h = _handles.get(device)
    if h is None:
        h = _handles[device] = Handle(device)
    return h


def _ptr(a: np.ndarray):
    return _vp(a.ctypes.data)


def _img(img) -> np.ndarray:
    a = np.ascontiguousarray(img)
    if a.dtype != np.uint8:
        raise TypeError(f"image must be uint8, got {a.dtype}")
    if a.ndim != 2:
        raise ValueError(f"image must be 2-d, got shape {a.shape}")
    return a


# ----------------------------------------------------------------------------------------------
# brute force
# ----------------------------------------------------------------------------------------------
def bf_encode(img, range_size=8, domain_size=16, strides=(1, 1), mode=BF_MODE_JIT, range_begin=0, range_end=None,
              flags=0, handle: Handle | None = None):
    """Host arrays in, host arrays out: (idx[n,4] int32, so[n,2] float64, R[n] float64)."""
    h = handle or default_handle()
    a = _img(img)
    H, W = a.shape
    if range_size <= 0 or H % range_size or W % range_size:
        raise ValueError(f"image {H}x{W} is not a multiple of range_size {range_size}")
    total = (H // range_size) * (W // range_size)
    if range_end is None:
        range_end = total
    n = max(0, range_end - range_begin)
    idx = np.zeros((n, 4), np.int32)
    so = np.zeros((n, 2), np.float64)
    R = np.zeros(n, np.float64)
    _check(lib().fic_bf_encode(h.ptr, _ptr(a), H, W, int(range_size), int(domain_size), int(strides[0]), int(strides[1]),
                               int(mode), int(range_begin), int(range_end), _ptr(idx), _ptr(so), _ptr(R),
                               int(flags) & ~DEVICE_PTRS, None), "fic_bf_encode")
    return idx, so, R


def bf_encode_dev(h: Handle, img_ptr: int, H: int, W: int, idx_ptr: int, so_ptr: int, R_ptr: int, range_size=8,
                  domain_size=16, strides=(1, 1), mode=BF_MODE_JIT, range_begin=0, range_end=None, flags=0, stream=0):
    """Device pointers in and out (no copies, no synchronisation)."""
    if range_end is None:
        range_end = (H // range_size) * (W // range_size)
    _check(lib().fic_bf_encode(h.ptr, _vp(img_ptr), H, W, int(range_size), int(domain_size), int(strides[0]),
                               int(strides[1]), int(mode), int(range_begin), int(range_end), _vp(idx_ptr), _vp(so_ptr),
                               _vp(R_ptr) if R_ptr else None, int(flags) | DEVICE_PTRS, _vp(stream) if stream else None),
           "fic_bf_encode")


def bf_encode_static_dev(img_ptr: int, H: int, W: int, idx_ptr: int, so_ptr: int, R_ptr: int, strides=(1, 1),
                         mode=BF_MODE_JIT, range_begin=0, range_end=None, stream=0):
    """`fic_bf_encode_static`: the exact 8x8 encode for callers that must not allocate device memory.

    Takes raw device pointers only (image, idx[n,4] int32, so[n,2] float64, R[n] float64 on the current CUDA
    device), needs no Handle, allocates nothing and does not synchronise.  Images up to 256x256; larger ones raise
    ValueError.  All calls on a GPU share one built-in scratch array, so do not overlap two of them in time, and
    make the first call on a device before any stream capture (it resolves the scratch address)."""
    if range_end is None:
        range_end = (H // 8) * (W // 8)
    _check(lib().fic_bf_encode_static(_vp(img_ptr), int(H), int(W), int(strides[0]), int(strides[1]), int(mode),
                                      int(range_begin), int(range_end), _vp(idx_ptr), _vp(so_ptr), _vp(R_ptr),
                                      _vp(stream) if stream else None), "fic_bf_encode_static")


def bf_screen_probe(img, range_begin=0, n_ranges=32, handle: Handle | None = None) -> np.ndarray:
    """Raw tensor-core correlation values u[position, range*8 + isometry] for up to 32 ranges (diagnostic)."""
    h = handle or default_handle()
    a = _img(img)
    H, W = a.shape
    NP = (H - 16) * (W - 16)
    rows = (NP + 127) // 128 * 128
    out = np.zeros((rows, 256), np.float32)
    _check(lib().fic_bf_screen_probe(h.ptr, _ptr(a), H, W, int(range_begin), int(n_ranges), _ptr(out), 0, None),
           "fic_bf_screen_probe")
    return out[:NP, :n_ranges * 8]


def bf_encode_batch(imgs, range_size=8, domain_size=16, mode=BF_MODE_JIT, flags=0, handle: Handle | None = None):
    h = handle or default_handle()
    a = np.ascontiguousarray(imgs)
    if a.dtype != np.uint8 or a.ndim != 3:
        raise TypeError("imgs must be uint8 [B,H,W]")
    B, H, W = a.shape
    n = (H // range_size) * (W // range_size)
    idx = np.zeros((B, n, 4), np.int32)
    so = np.zeros((B, n, 2), np.float64)
    R = np.zeros((B, n), np.float64)
    _check(lib().fic_bf_encode_batch(h.ptr, _ptr(a), B, H, W, int(range_size), int(domain_size), int(mode), _ptr(idx),
                                     _ptr(so), _ptr(R), int(flags) & ~DEVICE_PTRS, None), "fic_bf_encode_batch")
    return idx, so, R


def bf_decode(idx, so, width, height, iterations=32, handle: Handle | None = None) -> np.ndarray:
    h = handle or default_handle()
    idx = np.ascontiguousarray(idx, np.int32).reshape(-1, 4)
    so = np.ascontiguousarray(so, np.float64).reshape(-1, 2)
    if idx.shape[0] != so.shape[0]:
        raise ValueError("idx and so disagree on the number of transforms")
    out = np.zeros((int(height), int(width)), np.uint8)
    _check(lib().fic_bf_decode(h.ptr, _ptr(idx), _ptr(so), idx.shape[0], int(width), int(height), int(iterations),
                               _ptr(out), 0, None), "fic_bf_decode")
    return out


def bf_decode_batch(idx, so, width, height, iterations=32, handle: Handle | None = None) -> np.ndarray:
    """idx [planes,n,4], so [planes,n,2] -> uint8 [planes,height,width]; all planes decode concurrently (one CTA each)."""
    h = handle or default_handle()
    idx = np.ascontiguousarray(idx, np.int32)
    so = np.ascontiguousarray(so, np.float64)
    if idx.ndim != 3 or idx.shape[2] != 4 or so.shape != idx.shape[:2] + (2,):
        raise ValueError("idx must be [planes,n,4] and so [planes,n,2]")
    out = np.zeros((idx.shape[0], int(height), int(width)), np.uint8)
    if idx.shape[0] == 0:
        return out
    _check(lib().fic_bf_decode_batch(h.ptr, idx.shape[0], _ptr(idx), _ptr(so), idx.shape[1], int(width), int(height),
                                     int(iterations), _ptr(out), 0, None), "fic_bf_decode_batch")
    return out


def average_subsample(arr, jit=False, handle: Handle | None = None) -> np.ndarray:
    h = handle or default_handle()
    a = np.ascontiguousarray(arr, np.float64)
    if a.ndim != 2:
        raise ValueError("arr must be 2-d")
    out = np.zeros((a.shape[0] // 2, a.shape[1] // 2), np.float64)
    _check(lib().fic_average_subsample(h.ptr, _ptr(a), a.shape[0], a.shape[1], int(bool(jit)), _ptr(out), 0, None),
           "fic_average_subsample")
    return out


# ----------------------------------------------------------------------------------------------
# quadtree / HV
# ----------------------------------------------------------------------------------------------
def qt_classify(img, blocks, handle: Handle | None = None) -> np.ndarray:
    """blocks[k] = (start_i, start_j, side, negate) -> out[k] = (mean_class, variance_class, rotation, orientation)."""
    h = handle or default_handle()
    a = _img(img)
    b = np.ascontiguousarray(blocks, np.int32).reshape(-1, 4)
    out = np.zeros_like(b)
    _check(lib().fic_qt_classify(h.ptr, _ptr(a), a.shape[0], a.shape[1], _ptr(b), b.shape[0], _ptr(out), 0, None),
           "fic_qt_classify")
    return out


def qt_encode(img, scale_bits, offset_bits, max_scale, full_class_search, only_positive_scale, min_range_side,
              max_range_side, error_tolerance, domain_hstep=-1, domain_vstep=-1, handle: Handle | None = None,
              img_ptr: int = 0, shape=None):
    """Returns (nodes[k,8] int32, vals[k,3] float64, n_roots): the preorder stream of include/fic.h.
    With img_ptr/shape the image is taken from device memory instead of `img`."""
    h = handle or default_handle()
    flags = 0
    if img_ptr:
        H, W = shape
        p = _vp(img_ptr)
        flags = DEVICE_PTRS
    else:
        a = _img(img)
        H, W = a.shape
        p = _ptr(a)
    cap = 2 * H * W + 64
    nodes = np.empty((cap, 8), np.int32)     # written by the library up to the returned count
    vals = np.empty((cap, 3), np.float64)
    nroots = C.c_int32(0)
    n = _check(lib().fic_qt_encode(h.ptr, p, H, W, int(scale_bits), int(offset_bits), float(max_scale),
                                   int(bool(full_class_search)), int(bool(only_positive_scale)), int(min_range_side),
                                   int(max_range_side), float(error_tolerance), int(domain_hstep), int(domain_vstep),
                                   _ptr(nodes), _ptr(vals), cap, C.byref(nroots), flags, None), "fic_qt_encode")
    return nodes[:n].copy(), vals[:n].copy(), int(nroots.value)


def hv_encode(img, min_range_height, min_range_width, max_range_height, max_range_width, scale_bits, offset_bits,
              err_threshold, max_scale=1.0, handle: Handle | None = None, img_ptr: int = 0, shape=None):
    h = handle or default_handle()
    flags = 0
    if img_ptr:
        H, W = shape
        p = _vp(img_ptr)
        flags = DEVICE_PTRS
    else:
        a = _img(img)
        H, W = a.shape
        p = _ptr(a)
    cap = 2 * H * W + 64
    nodes = np.empty((cap, 8), np.int32)
    vals = np.empty((cap, 3), np.float64)
    n = _check(lib().fic_hv_encode(h.ptr, p, H, W, int(min_range_height), int(min_range_width), int(max_range_height),
                                   int(max_range_width), int(scale_bits), int(offset_bits), float(err_threshold),
                                   float(max_scale), _ptr(nodes), _ptr(vals), cap, flags, None), "fic_hv_encode")
    return nodes[:n].copy(), vals[:n].copy()


def hv_search(img, rects, scale_bits, offset_bits, max_scale=1.0, handle: Handle | None = None):
    """rects[k] = (i, j, w, h) -> (rec[k,6] = di,dj,orient,rot,qs,qo ; vals[k,3] = s,o,err)."""
    h = handle or default_handle()
    a = _img(img)
    r = np.ascontiguousarray(rects, np.int32).reshape(-1, 4)
    rec = np.zeros((r.shape[0], 6), np.int32)
    vals = np.zeros((r.shape[0], 3), np.float64)
    _check(lib().fic_hv_search(h.ptr, _ptr(a), a.shape[0], a.shape[1], _ptr(r), r.shape[0], int(scale_bits),
                               int(offset_bits), float(max_scale), _ptr(rec), _ptr(vals), 0, None), "fic_hv_search")
    return rec, vals


def tree_decode(kind, leaves, so, H, W, iterations=ITER_AUTO, stop_on_relative_error=5e-3, min_iterations=2,
                max_iterations=10, f32_affine=False, handle: Handle | None = None, deblock=False, out_f64=False):
    """leaves[k] = (range_i, range_j, height, width, start_i, start_j, orientation, rotation); so[k] = (scale, offset).
    Returns (uint8 image, sweeps run).  deblock (quadtree): the reference's deblocking post-processor runs on the
    float image before the clip; out_f64: the float64 image is returned unclipped instead of the bytes."""
    h = handle or default_handle()
    L = np.ascontiguousarray(leaves, np.int32).reshape(-1, 8)
    s = np.ascontiguousarray(so, np.float64).reshape(-1, 2)
    if L.shape[0] != s.shape[0]:
        raise ValueError("leaves and so disagree on the number of leaves")
    out = np.zeros((int(H), int(W)), np.float64 if out_f64 else np.uint8)
    done = C.c_int32(0)
    flags = (QT_DEBLOCK if deblock else 0) | (TREE_OUT_F64 if out_f64 else 0)
    _check(lib().fic_tree_decode(h.ptr, int(kind), _ptr(L), _ptr(s), L.shape[0], int(H), int(W), int(iterations),
                                 float(stop_on_relative_error), int(min_iterations), int(max_iterations),
                                 int(bool(f32_affine)), _ptr(out), C.byref(done), flags, None), "fic_tree_decode")
    return out, int(done.value)


def tree_decode_dev(h: Handle, kind, leaves_ptr: int, so_ptr: int, n: int, H: int, W: int, out_ptr: int,
                    iterations=ITER_AUTO, stop_on_relative_error=5e-3, min_iterations=2, max_iterations=10,
                    f32_affine=False, flags=0, stream=0) -> int:
    """fic_tree_decode on device buffers (leaves int32 [n,8], so float64 [n,2], out uint8 [H,W] or float64 with
    TREE_OUT_F64); returns the sweeps run.  The AUTO stop rule still reads its two norms back every sweep."""
    done = C.c_int32(0)
    _check(lib().fic_tree_decode(h.ptr, int(kind), _vp(leaves_ptr), _vp(so_ptr), int(n), int(H), int(W), int(iterations),
                                 float(stop_on_relative_error), int(min_iterations), int(max_iterations),
                                 int(bool(f32_affine)), _vp(out_ptr), C.byref(done), DEVICE_PTRS | flags,
                                 _vp(stream) if stream else None),
           "fic_tree_decode")
    return int(done.value)


def qt_deblock(img, leaves, weights=None, max_depth_range_size=-1, handle: Handle | None = None) -> np.ndarray:
    """DeblockingQuadtreePostprocessor.postprocess on a float64 image; leaves as for tree_decode, in any order."""
    h = handle or default_handle()
    a = np.ascontiguousarray(img, np.float64)
    if a.ndim != 2:
        raise ValueError("expected a 2-D image")
    L = np.ascontiguousarray(leaves, np.int32).reshape(-1, 8)
    w = None if weights is None else np.ascontiguousarray(weights, np.float64).reshape(4)
    out = np.empty_like(a)
    _check(lib().fic_qt_deblock(h.ptr, _ptr(a), _ptr(L), L.shape[0], a.shape[0], a.shape[1], None if w is None else _ptr(w),
                                int(max_depth_range_size), _ptr(out), 0, None), "fic_qt_deblock")
    return out


# ----------------------------------------------------------------------------------------------
# colour planes
# ----------------------------------------------------------------------------------------------
def _subsample_flag(chroma_subsampling) -> int:
    if chroma_subsampling in ("4:4:4", 0, False):
        return 0
    if chroma_subsampling in ("4:2:0", 1, True):
        return 1
    raise Exception("Unsupported chroma subsampling")


def rgb_to_ycbcr(rgb, chroma_subsampling="4:2:0", handle: Handle | None = None):
    """uint8 [N,M,3] -> (y, cb, cr) uint8 planes (utils.rgb_to_ycbcr)."""
    h = handle or default_handle()
    a = np.ascontiguousarray(rgb, np.uint8)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected an [N, M, 3] uint8 image")
    sub = _subsample_flag(chroma_subsampling)
    N, M = a.shape[:2]
    cs = (N // 2, (M + 1) // 2) if sub else (N, M)
    y, cb, cr = np.empty((N, M), np.uint8), np.empty(cs, np.uint8), np.empty(cs, np.uint8)
    _check(lib().fic_rgb_to_ycbcr(h.ptr, _ptr(a), N, M, sub, _ptr(y), _ptr(cb), _ptr(cr), 0, None), "fic_rgb_to_ycbcr")
    return y, cb, cr


def rgb_to_ycbcr_dev(h: Handle, rgb_ptr: int, N: int, M: int, chroma_subsampling, y_ptr: int, cb_ptr: int, cr_ptr: int):
    _check(lib().fic_rgb_to_ycbcr(h.ptr, _vp(rgb_ptr), int(N), int(M), _subsample_flag(chroma_subsampling), _vp(y_ptr),
                                  _vp(cb_ptr), _vp(cr_ptr), DEVICE_PTRS, None), "fic_rgb_to_ycbcr")


def ycbcr_to_rgb_dev(h: Handle, y_ptr: int, cb_ptr: int, cr_ptr: int, N: int, M: int, chroma_subsampling, rgb_ptr: int):
    _check(lib().fic_ycbcr_to_rgb(h.ptr, _vp(y_ptr), _vp(cb_ptr), _vp(cr_ptr), int(N), int(M),
                                  _subsample_flag(chroma_subsampling), _vp(rgb_ptr), DEVICE_PTRS, None), "fic_ycbcr_to_rgb")


def ycbcr_to_rgb(y, cb, cr, chroma_subsampling="4:2:0", handle: Handle | None = None) -> np.ndarray:
    """(y, cb, cr) uint8 planes -> uint8 [N,M,3] (utils.ycbcr_to_rgb)."""
    h = handle or default_handle()
    y, cb, cr = (np.ascontiguousarray(p, np.uint8) for p in (y, cb, cr))
    sub = _subsample_flag(chroma_subsampling)
    N, M = y.shape
    cs = (N // 2, (M + 1) // 2) if sub else (N, M)
    if cb.shape != cs or cr.shape != cs:
        raise ValueError(f"chroma planes {cb.shape}, {cr.shape} do not match a {N}x{M} image at this subsampling")
    out = np.empty((N, M, 3), np.uint8)
    _check(lib().fic_ycbcr_to_rgb(h.ptr, _ptr(y), _ptr(cb), _ptr(cr), N, M, sub, _ptr(out), 0, None), "fic_ycbcr_to_rgb")
    return out


# ----------------------------------------------------------------------------------------------
# container files (host only)
# ----------------------------------------------------------------------------------------------
def tree_read_nodes(kind, data: bytes, bit_pos: int, root_exps, raw_width, huff_root, huff_table, cap: int):
    """Node section of a `.fqt` / HV stream -> (preorder nodes[k,8] int32, first bit after the section)."""
    buf = np.frombuffer(data, np.uint8)
    exps = np.ascontiguousarray(root_exps if root_exps is not None else [0], np.int32)
    rw = np.ascontiguousarray(raw_width, np.int32).reshape(4)
    hr = np.ascontiguousarray(huff_root, np.int32).reshape(4)
    ht = np.ascontiguousarray(huff_table, np.int32).reshape(-1, 3)
    rows = ht.shape[0]
    if rows == 0:
        ht = np.zeros((1, 3), np.int32)
    nodes = np.empty((int(cap), 8), np.int32)
    end = C.c_int64(0)
    n = _check(lib().fic_tree_read_nodes(int(kind), buf.ctypes.data_as(_vp), buf.size, int(bit_pos), _ptr(exps),
                                         exps.shape[0] if kind == TREE_QUADTREE else 1, _ptr(rw), _ptr(hr), _ptr(ht),
                                         rows, _ptr(nodes), int(cap), C.byref(end)), "fic_tree_read_nodes")
    return nodes[:n].copy(), int(end.value)


def huffman_read_tree(data: bytes, bit_pos: int, val_bits: int):
    """One serialised Huffman tree -> (table[m,3] = left, right, value ; first bit after it)."""
    buf = np.frombuffer(data, np.uint8)
    cap = 2 * (1 << min(int(val_bits), 20)) + 2          # a tree over val_bits-bit symbols has at most 2^(bits+1) - 1 nodes
    table = np.empty((cap, 3), np.int32)
    end = C.c_int64(0)
    n = _check(lib().fic_huffman_read_tree(buf.ctypes.data_as(_vp), buf.size, int(bit_pos), int(val_bits), _ptr(table), cap,
                                           C.byref(end)), "fic_huffman_read_tree")
    return table[:n].copy(), int(end.value)


def pack_fields(values, widths, pending: int = 0, pending_bits: int = 0):
    """BitBuffer.write for whole columns: returns (bytes of the completed 32-bit words, pending value, pending bit count)."""
    v = np.ascontiguousarray(values, np.uint64).ravel()
    w = np.ascontiguousarray(widths, np.int64).ravel()
    if v.size != w.size:
        raise ValueError("values and widths differ in length")
    out = np.empty(((int(w.sum()) + int(pending_bits)) // 32 + 1) * 4, np.uint8)
    rest, rest_bits = C.c_uint32(0), C.c_int(0)
    n = _check(lib().fic_pack_fields(_ptr(v), _ptr(w), v.size, int(pending) & 0xffffffff, int(pending_bits), _ptr(out), out.size,
                                     C.byref(rest), C.byref(rest_bits)), "fic_pack_fields")
    return out[:n].tobytes(), int(rest.value), int(rest_bits.value)


def hv_layout(nodes, H: int, W: int):
    """(rects[k,4] = i, j, w, h ; depth[k]) of every node of an HV preorder stream."""
    a = np.ascontiguousarray(nodes, np.int32).reshape(-1, 8)
    rects = np.empty((a.shape[0], 4), np.int32)
    depth = np.empty(a.shape[0], np.int32)
    _check(lib().fic_hv_layout(_ptr(a), a.shape[0], int(H), int(W), _ptr(rects), _ptr(depth)), "fic_hv_layout")
    return rects, depth
