"""Generates tests/golden/*.npz from the REFERENCE ITSELF, run in the build container:
  * the reference's own CPU headers (/root/reference/include/cpp, compiled unmodified into
    oracle/_ref/libvip_ref_cpu.so by oracle/Makefile) on the reference's own test fixtures
    (std::mt19937(42) % 255, test/random_array.hpp:10-19; 50x50x3, test/bilateral_filter.cu:37-41), and
  * cv2.bilateralFilter, the known-answer oracle of the reference's test (test/bilateral_filter.cu:52).
/root/reference does not exist on the GPU box, so the vectors are committed.  Re-run:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle as O  # noqa: E402


def main():
    R = O.ref_cpu()
    R.ref_set_num_threads(4)
    w = h = 50
    src = O.mt_u8(w * h * 3).reshape(h, w, 3)           # test/bilateral_filter.cu:40
    out = {}
    out["src_mt42"] = src

    def run_bf(img, k, ss, sc):
        d = np.empty_like(img)
        R.ref_bilateral_filter(img.ctypes.data, d.ctypes.data, img.shape[1], img.shape[0], k, ss, sc)
        return d

    def run_jbf(img, g, k, ss, sc):
        d = np.empty_like(img)
        R.ref_joint_bilateral_filter(img.ctypes.data, g.ctypes.data, d.ctypes.data, img.shape[1], img.shape[0], k, ss, sc)
        return d

    def run_abf(img, k, ss, sc):
        d = np.empty_like(img)
        R.ref_adaptive_bilateral_filter(img.ctypes.data, d.ctypes.data, img.shape[1], img.shape[0], k, ss, sc)
        return d

    out["bf_k9"] = run_bf(src, 9, 10.0, 30.0)            # CppBilateralFilter, test/bilateral_filter.cu:35-63
    out["bf_k11"] = run_bf(src, 11, 10.0, 30.0)
    out["jbf_k9_self"] = run_jbf(src, src, 9, 10.0, 30.0)  # CppJointBilateralFilter (guide == src in the reference test)
    guide = O.rand_image(w, h, seed=7)
    out["guide_seed7"] = guide
    out["jbf_k9_guide7"] = run_jbf(src, guide, 9, 10.0, 30.0)
    out["abf_k9"] = run_abf(src, 9, 10.0, 30.0)          # test/adaptive_bilateral_filter.cu:139-164
    out["abf_k15"] = run_abf(src, 15, 10.0, 30.0)
    smooth = O.smooth_image(64, 48, seed=5)
    out["smooth_64x48"] = smooth
    out["bf_k11_smooth"] = run_bf(smooth, 11, 10.0, 30.0)
    out["abf_k15_smooth"] = run_abf(smooth, 15, 10.0, 30.0)

    # gradient, test/gradient.cu:38-143 (u8 / f32 x 1 / 3 channels)
    for ch in (1, 3):
        s8 = O.mt_u8(w * h * ch).reshape(h, w, ch)
        g = np.empty((h, w), np.float32)
        assert R.ref_gradient(s8.ctypes.data, g.ctypes.data, w, h, ch, 0) == 0
        out[f"grad_u8_c{ch}"] = g
        sf = O.mt_f32(w * h * ch).reshape(h, w, ch)
        g = np.empty((h, w), np.float32)
        assert R.ref_gradient(sf.ctypes.data, g.ctypes.data, w, h, ch, 1) == 0
        out[f"grad_f32_c{ch}"] = g

    # BTF stages, test/bilateral_texture_filter.cu:138-206
    mag = O.mt_f32(w * h).reshape(h, w)
    bl = np.empty((h, w, 3), np.float32)
    rt = np.empty((h, w), np.float32)
    R.ref_compute_blur_and_rtv(src.ctypes.data, mag.ctypes.data, bl.ctypes.data, rt.ctypes.data, w, h, 9)
    out["btf_blurred_cpu"] = bl
    out["btf_rtv_cpu"] = rt
    blurred_in = O.mt_f32(w * h * 3).reshape(h, w, 3)
    rtv_in = O.mt_f32(w * h, 1.0).reshape(h, w)
    gd = np.empty((h, w, 3), np.uint8)
    R.ref_compute_guide(blurred_in.ctypes.data, rtv_in.ctypes.data, gd.ctypes.data, w, h, 9)
    out["btf_guide_cpu"] = gd
    d = np.empty_like(smooth)
    R.ref_bilateral_texture_filter(smooth.ctypes.data, d.ctypes.data, 64, 48, 9, 3)
    out["btf_k9_n3_smooth_cpu"] = d

    # integral image: every rectangle of a 5x5 image, r in {1,3,5} (test/border_replicated_integral_image.cpp:19-57)
    small = O.mt_u8(5 * 5 * 3).reshape(5, 5, 3)
    out["ii_src_5x5"] = small
    for r in (1, 3, 5):
        rects = [(x0, y0, x1, y1) for y0 in range(-r, 5 + r) for x0 in range(-r, 5 + r)
                 for y1 in range(y0, 5 + r) for x1 in range(x0, 5 + r)]
        rects = np.asarray(rects, np.int32)
        res = np.empty((len(rects), 3), np.int32)
        R.ref_integral_u8c3_get(small.ctypes.data, 5, 5, r, rects.ctypes.data, len(rects), res.ctypes.data)
        out[f"ii_rects_r{r}"] = rects
        out[f"ii_sums_r{r}"] = res

    # third-party known answer of the reference's own test
    try:
        import cv2
        out["cv2_bilateral_k9"] = cv2.bilateralFilter(src, 9, 30, 10, borderType=cv2.BORDER_REPLICATE)
        out["cv2_version"] = np.frombuffer(cv2.__version__.encode(), np.uint8)
    except ImportError:
        print("cv2 missing: cv2_bilateral_k9 not regenerated")
    np.savez_compressed(os.path.join(HERE, "reference_vectors.npz"), **out)
    for k2 in sorted(out):
        print(f"{k2:28s} {out[k2].shape} {out[k2].dtype} fnv1a64={O.fnv1a64(np.ascontiguousarray(out[k2]).tobytes())}")


if __name__ == "__main__":
    main()
