"""ctypes bindings for the checkers under oracle/ (TEST INFRASTRUCTURE ONLY).

Three libraries, all optional at import time (each accessor raises a clear error when missing):
  oracle/libvip_oracle.so          plain-C restatement (oracle/vip_oracle.c)
  oracle/_ref/libvip_ref_cpu.so    the reference's own CPU headers, compiled unmodified (needs /root/reference to build)
  oracle/_ref/libvip_ref_cuda.so   the reference's own CUDA kernels for sm_100a (GPU box only)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ARITH_CUDA, ARITH_CPU = 0, 1

_vp, _i, _f = C.c_void_p, C.c_int, C.c_float


def build_oracle():
    """Compile the C restatement (and oracle/_ref when /root/reference is present)."""
    subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "libvip_oracle.so"])
    if os.path.isdir("/root/reference/include/cpp"):
        subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "ref"])


def _load(path, what):
    if not os.path.exists(path):
        raise FileNotFoundError(f"{what} not built: {path} (run `make -C oracle`)")
    return C.CDLL(path)


_cache = {}


def oracle():
    if "o" not in _cache:
        p = os.path.join(ORACLE_DIR, "libvip_oracle.so")
        if not os.path.exists(p):
            build_oracle()
        L = _load(p, "oracle restatement")
        L.vipo_build_luts.argtypes = [_i, _f, _f, _i, _i, _vp, _vp]
        L.vipo_bilateral_filter.argtypes = [_vp, _vp, _i, _i, _i, _f, _f, _i]
        L.vipo_joint_bilateral_filter.argtypes = [_vp, _vp, _vp, _i, _i, _i, _f, _f, _i]
        L.vipo_joint_bilateral_filter_rows.argtypes = [_vp, _vp, _vp, _i, _i, _i, _f, _f, _i, _i, _i]
        L.vipo_adaptive_bilateral_filter.argtypes = [_vp, _vp, _i, _i, _i, _f, _f, _i]
        L.vipo_adaptive_bilateral_filter_rows.argtypes = [_vp, _vp, _i, _i, _i, _f, _f, _i, _i, _i]
        L.vipo_integral_u8c3.argtypes = [_vp, _i, _i, _i, _vp]
        L.vipo_integral_get.argtypes = [_vp, _i, _i, _i, _i, _i, _i, _vp]
        L.vipo_gradient_u8.argtypes = [_vp, _vp, _i, _i, _i]
        L.vipo_gradient_f32.argtypes = [_vp, _vp, _i, _i, _i, _i]
        L.vipo_blur_rtv.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _i, _i]
        L.vipo_guide.argtypes = [_vp, _vp, _vp, _i, _i, _i, _i]
        L.vipo_bilateral_texture_filter.argtypes = [_vp, _vp, _i, _i, _i, _i, _i]
        L.vipo_set_num_threads.argtypes = [_i]
        L.vipo_get_num_threads.restype = _i
        _cache["o"] = L
    return _cache["o"]


def have_ref_cpu():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libvip_ref_cpu.so"))


def have_ref_cuda():
    return os.path.exists(os.path.join(ORACLE_DIR, "_ref", "libvip_ref_cuda.so"))


def ref_cpu():
    if "rc" not in _cache:
        L = _load(os.path.join(ORACLE_DIR, "_ref", "libvip_ref_cpu.so"), "reference CPU build")
        L.ref_set_num_threads.argtypes = [_i]
        L.ref_get_num_threads.restype = _i
        L.ref_bilateral_filter.argtypes = [_vp, _vp, _i, _i, _i, _f, _f]
        L.ref_joint_bilateral_filter.argtypes = [_vp, _vp, _vp, _i, _i, _i, _f, _f]
        L.ref_adaptive_bilateral_filter.argtypes = [_vp, _vp, _i, _i, _i, _f, _f]
        L.ref_gradient.argtypes = [_vp, _vp, _i, _i, _i, _i]
        L.ref_gradient.restype = _i
        L.ref_compute_blur_and_rtv.argtypes = [_vp, _vp, _vp, _vp, _i, _i, _i]
        L.ref_compute_guide.argtypes = [_vp, _vp, _vp, _i, _i, _i]
        L.ref_bilateral_texture_filter.argtypes = [_vp, _vp, _i, _i, _i, _i]
        L.ref_integral_u8c3_get.argtypes = [_vp, _i, _i, _i, _vp, _i, _vp]
        L.ref_integral_u8c3_table.argtypes = [_vp, _i, _i, _i, _vp]
        _cache["rc"] = L
    return _cache["rc"]


def ref_cuda():
    if "rg" not in _cache:
        L = _load(os.path.join(ORACLE_DIR, "_ref", "libvip_ref_cuda.so"), "reference CUDA build")
        for name in ("refcu_bf_create", "refcu_abf_create"):
            getattr(L, name).argtypes = [_i, _i, _i, _f, _f]
            getattr(L, name).restype = _vp
        L.refcu_btf_create.argtypes = [_i, _i, _i, _i]
        L.refcu_btf_create.restype = _vp
        for name in ("refcu_bf_destroy", "refcu_abf_destroy", "refcu_btf_destroy"):
            getattr(L, name).argtypes = [_vp]
        L.refcu_bf_run.argtypes = [_vp, _vp, _vp]
        L.refcu_jbf_run.argtypes = [_vp, _vp, _vp, _vp]
        L.refcu_abf_run.argtypes = [_vp, _vp, _vp]
        L.refcu_btf_run.argtypes = [_vp, _vp, _vp]
        L.refcu_btf_blur_rtv.argtypes = [_vp, _vp, _vp, _vp, _vp, _i, _i]
        L.refcu_btf_guide.argtypes = [_vp, _vp, _vp, _vp, _i, _i]
        L.refcu_gradient_u8.argtypes = [_vp, _vp, _i, _i, _i]
        L.refcu_gradient_f32.argtypes = [_vp, _vp, _i, _i, _i]
        _cache["rg"] = L
    return _cache["rg"]


# ---------------------------------------------------------------- fixtures (reference: test/random_array.hpp)
def mt_u8(n, mx=255):
    """std::mt19937(42)() % mx as uint8 -- the reference's integer fixture (test/random_array.hpp:10-19)."""
    rs = np.random.RandomState(42)
    return (rs.randint(0, 2**32, size=n, dtype=np.uint64) % mx).astype(np.uint8)


def mt_f32(n, mx=255.0):
    """max * float(mt19937(42)()) / UINT32_MAX (test/random_array.hpp:22-31); float arithmetic as in C++."""
    rs = np.random.RandomState(42)
    r = rs.randint(0, 2**32, size=n, dtype=np.uint64).astype(np.float32)
    return (np.float32(mx) * r / np.float32(4294967295.0)).astype(np.float32)


def rand_image(w, h, seed=42, lo=0, hi=256):
    rs = np.random.RandomState(seed)
    return rs.randint(lo, hi, size=(h, w, 3)).astype(np.uint8)


def smooth_image(w, h, seed=7):
    """Natural-ish content: smooth gradients + edges + mild noise (small colour distances dominate)."""
    rs = np.random.RandomState(seed)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    img = np.stack([
        127 + 100 * np.sin(xx / 17.0) * np.cos(yy / 23.0),
        127 + 90 * np.sin((xx + yy) / 31.0),
        ((xx // 16 + yy // 16) % 2) * 180 + 30,
    ], axis=-1)
    img += rs.normal(0, 6, size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


# ---------------------------------------------------------------- numpy-facing helpers over the C oracle
def _p(a):
    return a.ctypes.data


def o_bf(src, ksize=9, ss=10.0, sc=30.0, arith=ARITH_CUDA):
    h, w, _ = src.shape
    src = np.ascontiguousarray(src)
    dst = np.empty_like(src)
    oracle().vipo_bilateral_filter(_p(src), _p(dst), w, h, ksize, ss, sc, arith)
    return dst


def o_jbf(src, guide, ksize=9, ss=10.0, sc=30.0, arith=ARITH_CUDA):
    h, w, _ = src.shape
    src = np.ascontiguousarray(src); guide = np.ascontiguousarray(guide)
    dst = np.empty_like(src)
    oracle().vipo_joint_bilateral_filter(_p(src), _p(guide), _p(dst), w, h, ksize, ss, sc, arith)
    return dst


def o_abf(src, ksize=9, ss=10.0, sc=30.0, arith=ARITH_CUDA):
    h, w, _ = src.shape
    src = np.ascontiguousarray(src)
    dst = np.empty_like(src)
    oracle().vipo_adaptive_bilateral_filter(_p(src), _p(dst), w, h, ksize, ss, sc, arith)
    return dst


def o_rows(fn, src, y0, rows, ksize, ss=10.0, sc=30.0, arith=ARITH_CUDA, guide=None):
    """Rows [y0, y0+rows) of a full-image filter (used for full-size spot checks)."""
    h, w, _ = src.shape
    dst = np.zeros_like(src)
    if fn == "abf":
        oracle().vipo_adaptive_bilateral_filter_rows(_p(src), _p(dst), w, h, ksize, ss, sc, arith, y0, rows)
    else:
        g = src if guide is None else guide
        oracle().vipo_joint_bilateral_filter_rows(_p(src), _p(g), _p(dst), w, h, ksize, ss, sc, arith, y0, rows)
    return dst[y0:y0 + rows]


def o_gradient(src, arith=ARITH_CUDA):
    src = np.ascontiguousarray(src)
    h, w = src.shape[:2]
    ch = 1 if src.ndim == 2 else src.shape[2]
    dst = np.empty((h, w), np.float32)
    if src.dtype == np.uint8:
        oracle().vipo_gradient_u8(_p(src), _p(dst), w, h, ch)
    else:
        oracle().vipo_gradient_f32(_p(src), _p(dst), w, h, ch, arith)
    return dst


def o_blur_rtv(image, magnitude, ksize=9, arith=ARITH_CUDA):
    h, w, _ = image.shape
    image = np.ascontiguousarray(image); magnitude = np.ascontiguousarray(magnitude, np.float32)
    blurred = np.empty((h, w, 3), np.float32)
    rtv = np.empty((h, w), np.float32)
    oracle().vipo_blur_rtv(_p(image), _p(magnitude), _p(blurred), _p(rtv), w, h, ksize, arith)
    return blurred, rtv


def o_guide(blurred, rtv, ksize=9, arith=ARITH_CUDA):
    h, w, _ = blurred.shape
    blurred = np.ascontiguousarray(blurred, np.float32); rtv = np.ascontiguousarray(rtv, np.float32)
    guide = np.empty((h, w, 3), np.uint8)
    oracle().vipo_guide(_p(blurred), _p(rtv), _p(guide), w, h, ksize, arith)
    return guide


def o_btf(src, ksize=9, nitr=3, arith=ARITH_CUDA):
    h, w, _ = src.shape
    src = np.ascontiguousarray(src)
    dst = np.empty_like(src)
    oracle().vipo_bilateral_texture_filter(_p(src), _p(dst), w, h, ksize, nitr, arith)
    return dst


def o_integral(src, radius):
    h, w, _ = src.shape
    src = np.ascontiguousarray(src)
    table = np.empty((h + 2 * radius + 1, w + 2 * radius + 1, 3), np.uint32)
    oracle().vipo_integral_u8c3(_p(src), w, h, radius, _p(table))
    return table


def fnv1a64(buf):
    hsh = 0xCBF29CE484222325
    for x in bytes(buf):
        hsh ^= x
        hsh = (hsh * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return "%016x" % hsh
